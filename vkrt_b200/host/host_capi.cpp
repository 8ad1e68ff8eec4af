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
// out: proj * view * scale(s), column-major: the matrix Rasterizer::update uploads (src/Rasterizer.cpp:189-191)
HOST_API void vkrt_host_camera_wvp(void* c, float s, float* out)
{
    const Camera* cam = static_cast<Camera*>(c);
    const vkrt::mat4 wvp = cam->getProjectionMatrix() * cam->getViewMatrix() * vkrt::scale(vkrt::vec3(s, s, s));
    for (int i = 0; i < 16; ++i)
    {
        out[i] = (&wvp.col[0].x)[i];
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

// ---------------------------------------------------------------------------------------------------------------
// Model access (glTF loader) for the Python tests / bench
// ---------------------------------------------------------------------------------------------------------------
#include "Model.hpp"

#include <cstring>

HOST_API void* vkrt_host_model_create(const char* modelsFolder, const char* filename)
{
    if (modelsFolder && modelsFolder[0])
    {
        Model::setModelsFolder(modelsFolder);
    }
    return new Model(filename);
}
HOST_API void vkrt_host_model_destroy(void* m)
{
    delete static_cast<Model*>(m);
}
HOST_API uint32_t vkrt_host_model_submesh_count(void* m)
{
    return static_cast<uint32_t>(static_cast<Model*>(m)->submeshes.size());
}
HOST_API void vkrt_host_model_submesh_info(void* m, uint32_t i, uint32_t* vertexCount, uint32_t* indexCount, int* material)
{
    const Model::Submesh& s = static_cast<Model*>(m)->submeshes[i];
    *vertexCount = static_cast<uint32_t>(s.vertices.size());
    *indexCount = static_cast<uint32_t>(s.indices.size());
    *material = s.material;
}
HOST_API void vkrt_host_model_submesh_copy(void* m, uint32_t i, void* vertices, void* indices)
{
    const Model::Submesh& s = static_cast<Model*>(m)->submeshes[i];
    std::memcpy(vertices, s.vertices.data(), s.vertices.size() * sizeof(Model::Vertex));
    std::memcpy(indices, s.indices.data(), s.indices.size() * sizeof(Model::Index));
}
HOST_API uint32_t vkrt_host_model_material_count(void* m)
{
    return static_cast<uint32_t>(static_cast<Model*>(m)->materials.size());
}
HOST_API void vkrt_host_model_material(void* m, uint32_t i, int* out3)
{
    const Model::Material& mat = static_cast<Model*>(m)->materials[i];
    out3[0] = mat.baseColor, out3[1] = mat.metallicRoughnessImage, out3[2] = mat.normalImage;
}
HOST_API uint32_t vkrt_host_model_image_count(void* m)
{
    return static_cast<uint32_t>(static_cast<Model*>(m)->images.size());
}
HOST_API void vkrt_host_model_image_info(void* m, uint32_t i, uint32_t* width, uint32_t* height, uint32_t* components, uint32_t* bits)
{
    const Model::Image& im = static_cast<Model*>(m)->images[i];
    *width = im.width, *height = im.height, *components = im.components, *bits = im.bitsPerChannel;
}
HOST_API void vkrt_host_model_image_copy(void* m, uint32_t i, void* dst)
{
    const Model::Image& im = static_cast<Model*>(m)->images[i];
    std::memcpy(dst, im.data.data(), im.data.size());
}
HOST_API uint64_t vkrt_host_model_buffer_sizes(void* m, uint64_t* indexBytes)
{
    *indexBytes = static_cast<Model*>(m)->indexBufferSizeInBytes;
    return static_cast<Model*>(m)->vertexBufferSizeInBytes;
}
// Decode one image file in memory (PNG / JPEG) to RGBA8; returns 0 on success.  dst may be NULL to query the size.
#include "ImageDecode.hpp"
HOST_API int vkrt_host_decode_image(const void* data, uint64_t size, uint32_t* width, uint32_t* height, void* dst, uint64_t dstBytes)
{
    vkrt::DecodedImage img;
    std::string err;
    if (!vkrt::decodeImage(static_cast<const uint8_t*>(data), static_cast<size_t>(size), img, err))
    {
        return 1;
    }
    *width = img.width, *height = img.height;
    if (dst && dstBytes >= img.rgba.size())
    {
        std::memcpy(dst, img.rgba.data(), img.rgba.size());
    }
    return 0;
}
HOST_API uint64_t vkrt_host_fnv1a(const void* data, uint64_t size)
{
    const unsigned char* p = static_cast<const unsigned char*>(data);
    uint64_t h = 1469598103934665603ull;
    for (uint64_t i = 0; i < size; ++i)
    {
        h ^= p[i];
        h *= 1099511628211ull;
    }
    return h;
}
