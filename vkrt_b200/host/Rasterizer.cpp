// Rasterizer.cpp — host driver of the rasteriser path; follows the reference's constructor sequence
// (src/Rasterizer.cpp:58-84) and frame loop (:100-196) with the Vulkan calls replaced by the C ABI.
#include "Rasterizer.hpp"
#include "Utils.hpp"

#include <algorithm>
#include <cstring>
#include <fstream>

Rasterizer::Rasterizer(Context& context) :
    m_context(context),
    m_device(context.getDevice()),
    m_camera(static_cast<float>(context.getWidth()) / static_cast<float>(context.getHeight())),
    m_lastRenderTime(std::chrono::steady_clock::now())
{
    loadModel();
    setupCamera();
    createColorImage();
    createTextures();
    createVertexAndIndexBuffer();
    updateTexturesDescriptorSets();
    m_model.reset();
}

Rasterizer::~Rasterizer()
{
    m_context.waitIdle();
    vkrt_context_synchronize(m_device);
    vkrt_frame_destroy(m_device, m_frame);
    vkrt_blas_destroy(m_device, m_geometry);
}

bool Rasterizer::render()
{
    const uint32_t imageIndex = m_context.acquireNextSwapchainImage();
    if (!update(imageIndex))
    {
        return false;
    }
    // the render pass of src/Rasterizer.cpp:119-147: clears, one indexed draw per submesh, 8x MSAA resolve
    VKRT_CHECK(m_device, vkrt_rasterize(m_device, m_geometry, m_wvp, VKRT_TRACE_DEFAULT, m_frame));
    m_context.submitCommandBuffers(m_frame);
    return true;
}

bool Rasterizer::update(uint32_t /*imageIndex*/)
{
    if (!m_context.update())
    {
        return false;
    }
    using namespace std::chrono;
    double deltaTime = m_context.getFixedDeltaTime();
    if (deltaTime <= 0.0)
    {
        deltaTime = static_cast<double>(duration_cast<nanoseconds>(steady_clock::now() - m_lastRenderTime).count()) / 1'000'000'000.0;
    }
    m_fps = static_cast<float>(1.0 / deltaTime);
    m_lastRenderTime = steady_clock::now();

    updateCamera(deltaTime);
    // src/Rasterizer.cpp:189-191
    const vkrt::mat4 scaleMatrix = vkrt::scale(vkrt::vec3(0.01f, 0.01f, 0.01f));
    const vkrt::mat4 wvpMatrix = m_camera.getProjectionMatrix() * m_camera.getViewMatrix() * scaleMatrix;
    std::memcpy(m_wvp, &wvpMatrix.col[0].x, sizeof(m_wvp));
    return true;
}

void Rasterizer::loadModel()
{
    m_model.reset(new Model("sponza/Sponza.gltf"));
}

void Rasterizer::setupCamera()
{
    // src/Rasterizer.cpp:207-211
    m_camera.setPosition({-4.0f, 2.0f, -0.2f});
    m_camera.setRotation({0.0f, 1.51f, 0.0f});
}

void Rasterizer::updateCamera(double deltaTime)
{
    // src/Rasterizer.cpp:213-264 (the same key map as the ray tracer)
    for (const Context::KeyEvent& keyEvent : m_context.getKeyEvents())
    {
        if (keyEvent.action == GLFW_PRESS || keyEvent.action == GLFW_REPEAT)
        {
            m_keysDown[keyEvent.key] = true;
        }
        if (keyEvent.action == GLFW_RELEASE)
        {
            m_keysDown[keyEvent.key] = false;
        }
    }
    const float translationAmout = 5.0f * static_cast<float>(deltaTime);
    const float rotationAmout = 1.5f * static_cast<float>(deltaTime);
    if (m_keysDown[GLFW_KEY_W]) m_camera.translate(m_camera.getForward() * translationAmout);
    if (m_keysDown[GLFW_KEY_S]) m_camera.translate(-m_camera.getForward() * translationAmout);
    if (m_keysDown[GLFW_KEY_A]) m_camera.translate(m_camera.getLeft() * translationAmout);
    if (m_keysDown[GLFW_KEY_D]) m_camera.translate(-m_camera.getLeft() * translationAmout);
    if (m_keysDown[GLFW_KEY_E]) m_camera.translate(m_camera.getUp() * translationAmout);
    if (m_keysDown[GLFW_KEY_Q]) m_camera.translate(-m_camera.getUp() * translationAmout);
    if (m_keysDown[GLFW_KEY_Z]) m_camera.rotate(c_up, rotationAmout);
    if (m_keysDown[GLFW_KEY_C]) m_camera.rotate(-c_up, rotationAmout);
}

void Rasterizer::createColorImage()
{
    // multisampled colour + depth attachments and the resolve target (src/Rasterizer.cpp:326-440) live inside the frame
    VKRT_CHECK(m_device, vkrt_frame_create(m_device, m_context.getWidth(), m_context.getHeight(), &m_frame));
}

void Rasterizer::createTextures()
{
    std::vector<VkrtImage> images(m_model->images.size());
    for (size_t i = 0; i < images.size(); ++i)
    {
        CHECK(m_model->images[i].components == 4 && m_model->images[i].bitsPerChannel == 8);
        images[i].width = m_model->images[i].width;
        images[i].height = m_model->images[i].height;
        images[i].rgba8 = m_model->images[i].data.data();
    }
    VKRT_CHECK(m_device, vkrt_textures_set(m_device, images.data(), ui32Size(images)));
    createMipmaps();
}

void Rasterizer::createMipmaps()
{
    // src/Rasterizer.cpp createMipmaps: the fragment shader's implicit level of detail samples this chain
    VKRT_CHECK(m_device, vkrt_textures_generate_mips(m_device));
}

void Rasterizer::createVertexAndIndexBuffer()
{
    // src/Rasterizer.cpp:1040-1085: all vertices, then all indices, per primitive {vertexCountOffset, firstIndex, indexCount};
    // vkCmdDrawIndexed adds vertexCountOffset to every index (:145), which is the rebased index buffer built here.
    std::vector<uint8_t> vertexData(m_model->vertexBufferSizeInBytes, 0);
    std::vector<Model::Index> indexData(m_model->indexBufferSizeInBytes / sizeof(Model::Index), 0);
    std::vector<VkrtGeometryRange> ranges;
    size_t indexCounter = 0, vertexByteOffset = 0;
    Model::Index vertexCountOffset = 0;
    uint64_t indexByteOffset = 0;
    for (const Model::Submesh& submesh : m_model->submeshes)
    {
        Model::Index highestIndex = 0;
        for (Model::Index index : submesh.indices)
        {
            indexData[indexCounter++] = vertexCountOffset + index;
            highestIndex = std::max(highestIndex, index);
        }
        ranges.push_back(VkrtGeometryRange{ui32Size(submesh.indices) / 3, highestIndex, indexByteOffset});
        vertexCountOffset += static_cast<Model::Index>(submesh.vertices.size());
        const size_t vertexSize = sizeof(Model::Vertex) * submesh.vertices.size();
        std::memcpy(&vertexData[vertexByteOffset], submesh.vertices.data(), vertexSize);
        vertexByteOffset += vertexSize;
        indexByteOffset += sizeof(Model::Index) * submesh.indices.size();
    }
    CHECK(indexCounter == indexData.size());
    VKRT_CHECK(m_device, vkrt_blas_build(m_device, vertexData.data(), vertexData.size(), sizeof(Model::Vertex), indexData.data(), indexData.size(),
                                         ranges.data(), ui32Size(ranges), VKRT_BUILD_PREFER_FAST_BUILD, &m_geometry));
    VKRT_CHECK(m_device, vkrt_blas_info(m_device, m_geometry, &m_geometryInfo));
}

void Rasterizer::updateTexturesDescriptorSets()
{
    // src/Rasterizer.cpp:143: every draw binds the texture set of its primitive's material; the fragment shader reads
    // only baseColor (shaders/shader.frag:15).  One row per submesh, as the ray tracer's material table.
    std::vector<VkrtSubmeshInfo> rows(m_model->submeshes.size());
    int indexBufferOffset = 0;
    for (size_t i = 0; i < rows.size(); ++i)
    {
        const Model::Submesh& submesh = m_model->submeshes[i];
        rows[i].baseColorTextureIndex = m_model->materials[submesh.material].baseColor;
        rows[i].metallicRoughnessTextureIndex = std::max(m_model->materials[submesh.material].metallicRoughnessImage, 0);
        rows[i].normalTextureIndex = std::max(m_model->materials[submesh.material].normalImage, 0);
        rows[i].indexBufferOffset = indexBufferOffset;
        indexBufferOffset += static_cast<int>(submesh.indices.size() / 3);
    }
    VKRT_CHECK(m_device, vkrt_materials_set(m_device, rows.data(), ui32Size(rows)));
}

void Rasterizer::readback(std::vector<uint8_t>& bgra8)
{
    bgra8.resize(static_cast<size_t>(m_context.getWidth()) * m_context.getHeight() * 4);
    const std::vector<void*>& images = m_context.getSwapchainImages();
    if (!images.empty())
    {
        m_context.waitIdle();
        memcpy(bgra8.data(), images[m_context.getLastPresentedImage()], bgra8.size());
        return;
    }
    VKRT_CHECK(m_device, vkrt_frame_read(m_device, m_frame, VKRT_AOV_BGRA8, bgra8.data(), 0));
}

bool Rasterizer::saveFrame(const std::string& path)
{
    std::vector<uint8_t> bgra;
    readback(bgra);
    std::ofstream f(path, std::ios::binary);
    if (!f)
    {
        return false;
    }
    f << "P6\n" << m_context.getWidth() << " " << m_context.getHeight() << "\n255\n";
    std::vector<uint8_t> rgb(bgra.size() / 4 * 3);
    for (size_t i = 0; i < bgra.size() / 4; ++i)
    {
        rgb[3 * i] = bgra[4 * i + 2], rgb[3 * i + 1] = bgra[4 * i + 1], rgb[3 * i + 2] = bgra[4 * i];
    }
    f.write(reinterpret_cast<const char*>(rgb.data()), static_cast<std::streamsize>(rgb.size()));
    return static_cast<bool>(f);
}

VkrtStats Rasterizer::getStats()
{
    VkrtStats stats{};
    VKRT_CHECK(m_device, vkrt_stats_get(m_device, &stats));
    return stats;
}
