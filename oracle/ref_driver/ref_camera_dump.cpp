// ref_camera_dump.cpp — drives the REFERENCE's own Camera (compiled in place from /root/reference/src/Camera.cpp with
// the vendored glm) and prints, for a list of poses, the matrices and the uniform-block fields exactly as
// Raytracer::update computes them (/root/reference/src/Raytracer.cpp:224-235).  Output: JSON with float bit patterns.
// Built only into oracle/_ref/ by oracle/build_ref.sh; used to generate tests/golden/camera_golden.json.
#include "Camera.hpp"
#include "Utils.hpp"

#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>

#include <cstdint>
#include <cstdio>
#include <cstring>

static void printBits(const char* name, const float* v, int n, bool last = false)
{
    printf("    \"%s\": [", name);
    for (int i = 0; i < n; ++i)
    {
        uint32_t u;
        std::memcpy(&u, &v[i], 4);
        printf("%u%s", u, i + 1 < n ? ", " : "");
    }
    printf("]%s\n", last ? "" : ",");
}

static void dump(const Camera& cam, const char* label, bool last)
{
    const glm::mat4 view = cam.getViewMatrix(), proj = cam.getProjectionMatrix();
    const glm::mat4 viewInverse = glm::inverse(view), projInverse = glm::inverse(proj);
    // Rasterizer::update, /root/reference/src/Rasterizer.cpp:189-190 (glm::scale(v) of gtx/transform.hpp is scale(mat4(1), v))
    const glm::mat4 scaleMatrix = glm::scale(glm::mat4(1.0f), glm::vec3(0.01f, 0.01f, 0.01f));
    const glm::mat4 wvpMatrix = proj * view * scaleMatrix;
    const glm::vec3 f = cam.getForward(), l = cam.getLeft(), u = cam.getUp(), p = cam.getPosition();
    printf("  {\n    \"label\": \"%s\",\n", label);
    printBits("position", &p.x, 3);
    printBits("forward", &f.x, 3);
    printBits("left", &l.x, 3);
    printBits("up", &u.x, 3);
    printBits("view", &view[0][0], 16);
    printBits("proj", &proj[0][0], 16);
    printBits("viewInverse", &viewInverse[0][0], 16);
    printBits("projInverse", &projInverse[0][0], 16);
    printBits("wvp", &wvpMatrix[0][0], 16, true);
    printf("  }%s\n", last ? "" : ",");
}

int main()
{
    printf("[\n");
    {
        Camera cam;
        dump(cam, "default", false);
    }
    {
        Camera cam; // Raytracer::setupCamera, src/Raytracer.cpp:267-271
        cam.setPosition({6.3f, 4.5f, -0.7f});
        cam.setRotation({0.0f, 1.57f, 0.0f});
        dump(cam, "raytracer_default", false);
        // config-2 style path with the reference's own input semantics (src/Raytracer.cpp:288-323), dt = 1/60
        const float dt = 1.0f / 60.0f;
        const float translationAmount = 5.0f * dt, rotationAmount = 1.5f * dt;
        for (int i = 0; i < 150; ++i)
        {
            cam.translate(cam.getForward() * translationAmount);
        }
        dump(cam, "after_150_W", false);
        for (int i = 0; i < 150; ++i)
        {
            cam.rotate(c_up, rotationAmount);
        }
        dump(cam, "after_150_W_150_Z", false);
    }
    {
        Camera cam; // Rasterizer::setupCamera, src/Rasterizer.cpp:207-211
        cam.setPosition({-4.0f, 2.0f, -0.2f});
        cam.setRotation({0.0f, 1.51f, 0.0f});
        dump(cam, "rasterizer_default", false);
    }
    {
        Camera cam;
        cam.setPosition({-25.0f, 45.0f, -20.0f});
        cam.setRotation({-0.6f, -2.1f, 0.1f});
        dump(cam, "pitch_yaw_roll", false);
    }
    {
        Camera cam;
        cam.setPosition({0.25f, -3.0f, 12.5f});
        cam.setRotation({0.3f, 0.0f, 0.0f});
        cam.translate({1.0f, 2.0f, 3.0f});
        cam.rotate(c_up, 0.7f);
        dump(cam, "translate_rotate", true);
    }
    printf("]\n");
    return 0;
}
