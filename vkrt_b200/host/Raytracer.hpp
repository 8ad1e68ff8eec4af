// Raytracer.hpp — same public interface as the reference's Raytracer (src/Raytracer.hpp:11-17): construct over a
// Context, call render() until it returns false.  The private steps keep the reference's names; they call the C ABI
// of libvkrt_b200 (include/vkrt.h) where the reference calls Vulkan.  Headless additions: readback() and saveFrame().
#pragma once

#include "Camera.hpp"
#include "Context.hpp"
#include "Model.hpp"

#include <chrono>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

class Raytracer final
{
public:
    Raytracer(Context& context);
    ~Raytracer();

    bool render();

    // Headless mode (BASELINE.json north_star: "a headless offscreen mode ... so frames can be read back")
    void setSamplesPerPixel(uint32_t spp) { m_spp = spp; }
    void setTextureLod(int mode); // VKRT_TEXTURE_LOD_ZERO (reference behaviour, default) or VKRT_TEXTURE_LOD_RAY_CONES
    void setInstances(const std::vector<VkrtInstance>& instances); // config 4; default = the reference's single 0.01-scale instance
    void readback(std::vector<uint8_t>& bgra8);                     // B8G8R8A8, tightly packed (root rank after the gather)
    void readbackAov(int aov, void* dst);
    bool saveFrame(const std::string& path);                        // binary PPM (P6), RGB order
    VkrtStats getStats();
    Camera& getCamera() { return m_camera; }
    const VkrtBlasInfo& getBlasInfo() const { return m_blasInfo; }

private:
    struct SubmeshIndexInfo
    {
        Model::Index maxVertex;
        uint32_t triangleCount;
        uint64_t indexByteOffset;
    };

    bool update(uint32_t imageIndex);

    void loadModel();
    void setupCamera();
    void updateCamera(double deltaTime);
    void createColorImage();
    void createTextures();
    void createMipmaps();
    void createVertexAndIndexBuffer();
    void createBLAS();
    void createTLAS();
    void updateMaterialIndexDescriptorSet();

    Context& m_context;
    VkrtContext* m_device;

    std::unique_ptr<Model> m_model{nullptr};
    Camera m_camera;
    std::chrono::steady_clock::time_point m_lastRenderTime;
    std::unordered_map<int, bool> m_keysDown;
    std::vector<SubmeshIndexInfo> m_submeshIndexInfos;
    std::vector<uint8_t> m_vertexData;
    std::vector<Model::Index> m_indexData;
    std::vector<VkrtInstance> m_instances;

    VkrtBlas* m_blas = nullptr;
    VkrtTlas* m_tlas = nullptr;
    VkrtFrame* m_frame = nullptr;
    VkrtBlasInfo m_blasInfo{};
    VkrtUniforms m_uniforms{};
    uint32_t m_spp = 1;
    uint32_t m_frameIndex = 0;
    float m_fps = 0.0f;
};
