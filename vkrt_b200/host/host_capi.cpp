// host_capi.cpp — C shim over the host C++ classes so the Python tests and bench.py drive the same
// Camera / uniform code as the headless C++ application.
#include "Camera.hpp"

#if defined(__GNUC__)
#define HOST_API extern "C" __attribute__((visibility("default")))
#else
#define HOST_API extern "C"
#endif

HOST_API void* vkrt_host_camera_create(float aspectRatio)
{
    return new Camera(aspectRatio);
}
HOST_API void vkrt_host_camera_destroy(void* c)
{
    delete static_cast<Camera*>(c);
}
HOST_API void vkrt_host_camera_set_position(void* c, float x, float y, float z)
{
    static_cast<Camera*>(c)->setPosition(vkrt::vec3(x, y, z));
}
HOST_API void vkrt_host_camera_set_rotation(void* c, float x, float y, float z)
{
    static_cast<Camera*>(c)->setRotation(vkrt::vec3(x, y, z));
}
HOST_API void vkrt_host_camera_translate(void* c, float x, float y, float z)
{
    static_cast<Camera*>(c)->translate(vkrt::vec3(x, y, z));
}
HOST_API void vkrt_host_camera_rotate(void* c, float ax, float ay, float az, float amount)
{
    static_cast<Camera*>(c)->rotate(vkrt::vec3(ax, ay, az), amount);
}
HOST_API void vkrt_host_camera_set_aspect(void* c, float aspectRatio)
{
    static_cast<Camera*>(c)->setAspectRatio(aspectRatio);
}
// out: position(3) forward(3) left(3) up(3)
HOST_API void vkrt_host_camera_axes(void* c, float* out)
{
    const Camera* cam = static_cast<Camera*>(c);
    const vkrt::vec3 v[4] = {cam->getPosition(), cam->getForward(), cam->getLeft(), cam->getUp()};
    for (int i = 0; i < 4; ++i)
    {
        out[3 * i] = v[i].x, out[3 * i + 1] = v[i].y, out[3 * i + 2] = v[i].z;
    }
}
// out: view(16) proj(16), column-major
HOST_API void vkrt_host_camera_matrices(void* c, float* out)
{
    const Camera* cam = static_cast<Camera*>(c);
    const vkrt::mat4& v = cam->getViewMatrix();
    const vkrt::mat4& p = cam->getProjectionMatrix();
    for (int i = 0; i < 16; ++i)
    {
        out[i] = (&v.col[0].x)[i];
        out[16 + i] = (&p.col[0].x)[i];
    }
}
HOST_API void vkrt_host_fill_uniforms(void* c, const float* lights16, VkrtUniforms* out)
{
    float lights[4][4];
    for (int i = 0; i < 16; ++i)
    {
        lights[i / 4][i % 4] = lights16 ? lights16[i] : c_lightPositions[i / 4][i % 4];
    }
    fillUniforms(*static_cast<Camera*>(c), lights, out);
}
