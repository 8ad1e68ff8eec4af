// Rasterizer.hpp — same public interface as the reference's Rasterizer (src/Rasterizer.hpp:11-17): construct over a
// Context, call render() until it returns false.  The private steps keep the reference's names; they call the C ABI of
// libvkrt_b200 (vkrt_rasterize and the scene set-up calls) where the reference records a Vulkan render pass.
// Not mirrored: the ImGui overlay (src/Rasterizer.cpp:149-159) and the window.
#pragma once

#include "Camera.hpp"
#include "Context.hpp"
#include "Model.hpp"

#include <chrono>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

class Rasterizer final
{
public:
    Rasterizer(Context& context);
    ~Rasterizer();

    bool render();

    // Headless mode
    void readback(std::vector<uint8_t>& bgra8); // B8G8R8A8, tightly packed
    bool saveFrame(const std::string& path);    // binary PPM (P6), RGB order
    VkrtStats getStats();
    Camera& getCamera() { return m_camera; }
    const VkrtBlasInfo& getBlasInfo() const { return m_geometryInfo; }

private:
    bool update(uint32_t imageIndex);

    void loadModel();
    void setupCamera();
    void updateCamera(double deltaTime);
    void createColorImage();
    void createTextures();
    void createMipmaps();
    void createVertexAndIndexBuffer();
    void updateTexturesDescriptorSets();

    Context& m_context;
    VkrtContext* m_device;

    std::unique_ptr<Model> m_model{nullptr};
    Camera m_camera;
    std::chrono::steady_clock::time_point m_lastRenderTime;
    std::unordered_map<int, bool> m_keysDown;

    VkrtBlas* m_geometry = nullptr; // device copy of the vertex and index data (the handle type of the C ABI)
    VkrtFrame* m_frame = nullptr;
    VkrtBlasInfo m_geometryInfo{};
    float m_wvp[16] = {};
    float m_fps = 0.0f;
};
