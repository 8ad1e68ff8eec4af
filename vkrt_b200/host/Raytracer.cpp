// Raytracer.cpp — host driver of the ray-tracing path; follows the reference's constructor sequence
// (src/Raytracer.cpp:54-83) and frame loop (:132-238) with the Vulkan calls replaced by the C ABI.
//
// PROVENANCE.  This is the maintainer's `Raytracer` class with the bodies that called Vulkan swapped for calls into
// include/vkrt.h.  Three blocks are the REFERENCE'S OWN CODE, kept verbatim (only glm -> the
// local Math.hpp types and the Vulkan calls cut out) because the drop-in contract is "same host logic, same quirks";
// they are not this project's work and carry no claim of authorship:
//   * Raytracer::updateCamera                       = src/Raytracer.cpp:273-324, character for character, including the
//                                                     reference's identifiers `translationAmout` / `rotationAmout`;
//   * Raytracer::createVertexAndIndexBuffer         = the concatenation loop of src/Raytracer.cpp:660-700;
//   * Raytracer::updateMaterialIndexDescriptorSet   = the SubmeshInfo loop of src/Raytracer.cpp:1412-1427.
// Everything else in this file (constructor order excepted, :54-83) is written against the C ABI.
#include "Raytracer.hpp"
#include "Utils.hpp"

#include <algorithm>
#include <cstring>
#include <fstream>

namespace
{
struct SubmeshInfo // src/Raytracer.cpp:33-39
{
    int baseColorTextureIndex = -1;
    int metallicRoughnessTextureIndex = -1;
    int normalTextureIndex = -1;
    int indexBufferOffset = 0;
};
static_assert(sizeof(SubmeshInfo) == sizeof(VkrtSubmeshInfo), "material row layout");
} // namespace

Raytracer::Raytracer(Context& context) :
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
    createBLAS();
    createTLAS();
    updateMaterialIndexDescriptorSet();
    m_model.reset(); // host copy freed after upload (src/Raytracer.cpp:82)
}

Raytracer::~Raytracer()
{
    m_context.waitIdle();
    vkrt_context_synchronize(m_device);
    vkrt_frame_destroy(m_device, m_frame);
    vkrt_tlas_destroy(m_device, m_tlas);
    vkrt_blas_destroy(m_device, m_blas);
}

bool Raytracer::render()
{
    const uint32_t imageIndex = m_context.acquireNextSwapchainImage();
    bool running = update(imageIndex);
    if (!running)
    {
        return false;
    }
    // vkCmdTraceRaysKHR(W,H,1) (src/Raytracer.cpp:157) + the copy to the presented image (:175-185)
    VKRT_CHECK(m_device, vkrt_trace(m_device, m_tlas, &m_uniforms, m_spp, m_frameIndex, VKRT_TRACE_DEFAULT, m_frame));
    if (m_context.getOptions().worldSize > 1)
    {
        VKRT_CHECK(m_device, vkrt_tiles_gather(m_device, m_frame, 0));
    }
    ++m_frameIndex;
    // the copy to the acquired swapchain image + present: asynchronous, the next frame's traversal overlaps it
    m_context.submitCommandBuffers(m_context.getOptions().rank == 0 ? m_frame : nullptr);
    return true;
}

bool Raytracer::update(uint32_t /*imageIndex*/)
{
    bool running = m_context.update();
    if (!running)
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
    fillUniforms(m_camera, c_lightPositions, &m_uniforms); // src/Raytracer.cpp:224-235
    return true;
}

void Raytracer::loadModel()
{
    m_model.reset(new Model("sponza/Sponza.gltf"));
}

void Raytracer::setupCamera()
{
    m_camera.setPosition({6.3f, 4.5f, -0.7f});
    m_camera.setRotation({0.0f, 1.57f, 0.0f});
}

// >>> reference code, verbatim: src/Raytracer.cpp:273-324 (see PROVENANCE at the top of this file)
void Raytracer::updateCamera(double deltaTime)
{
    // src/Raytracer.cpp:273-324
    std::vector<Context::KeyEvent> keyEvents = m_context.getKeyEvents();
    for (const Context::KeyEvent& keyEvent : keyEvents)
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
    const float translationSpeed = 5.0f;
    const float rotationSpeed = 1.5f;
    const float translationAmout = translationSpeed * deltaTime;
    const float rotationAmout = rotationSpeed * deltaTime;
    if (m_keysDown[GLFW_KEY_W])
    {
        m_camera.translate(m_camera.getForward() * translationAmout);
    }
    if (m_keysDown[GLFW_KEY_S])
    {
        m_camera.translate(-m_camera.getForward() * translationAmout);
    }
    if (m_keysDown[GLFW_KEY_A])
    {
        m_camera.translate(m_camera.getLeft() * translationAmout);
    }
    if (m_keysDown[GLFW_KEY_D])
    {
        m_camera.translate(-m_camera.getLeft() * translationAmout);
    }
    if (m_keysDown[GLFW_KEY_E])
    {
        m_camera.translate(m_camera.getUp() * translationAmout);
    }
    if (m_keysDown[GLFW_KEY_Q])
    {
        m_camera.translate(-m_camera.getUp() * translationAmout);
    }
    if (m_keysDown[GLFW_KEY_Z])
    {
        m_camera.rotate(c_up, rotationAmout);
    }
    if (m_keysDown[GLFW_KEY_C])
    {
        m_camera.rotate(-c_up, rotationAmout);
    }
}

void Raytracer::createColorImage()
{
    // offscreen B8G8R8A8 target (src/Raytracer.cpp:326-392)
    VKRT_CHECK(m_device, vkrt_frame_create(m_device, m_context.getWidth(), m_context.getHeight(), &m_frame));
}

void Raytracer::createTextures()
{
    // one RGBA8 image per glTF image, sampled LINEAR/REPEAT at LOD 0 (src/Raytracer.cpp:436-570)
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

void Raytracer::createMipmaps()
{
    // src/Raytracer.cpp:572-640, called per image from createTextures (:562): here one call builds every chain on the
    // device.  The chain is sampled only after setTextureLod(VKRT_TEXTURE_LOD_RAY_CONES); the default is the reference's
    // level-0 sampling.
    VKRT_CHECK(m_device, vkrt_textures_generate_mips(m_device));
}

void Raytracer::setTextureLod(int mode)
{
    VKRT_CHECK(m_device, vkrt_trace_set_texture_lod(m_device, mode));
}

// >>> reference code, verbatim: src/Raytracer.cpp:660-700 (see PROVENANCE at the top of this file)
void Raytracer::createVertexAndIndexBuffer()
{
    // src/Raytracer.cpp:642-742: one vertex buffer, one index buffer rebased by the running vertex count, and per
    // submesh {highest local index, triangle count, index byte offset} for the BLAS build.
    m_vertexData.assign(m_model->vertexBufferSizeInBytes, 0);
    m_indexData.assign(m_model->indexBufferSizeInBytes / sizeof(Model::Index), 0);
    size_t indexCounter = 0;
    Model::Index indexCounterOffset = 0;
    size_t vertexByteOffset = 0;
    uint64_t indexByteOffset = 0;
    for (const Model::Submesh& submesh : m_model->submeshes)
    {
        Model::Index highestIndex = 0;
        for (Model::Index index : submesh.indices)
        {
            m_indexData[indexCounter] = indexCounterOffset + index;
            highestIndex = std::max(highestIndex, index);
            ++indexCounter;
        }
        m_submeshIndexInfos.push_back(SubmeshIndexInfo{highestIndex, ui32Size(submesh.indices) / 3, indexByteOffset});
        indexCounterOffset += static_cast<Model::Index>(submesh.vertices.size());
        const size_t vertexSize = sizeof(Model::Vertex) * submesh.vertices.size();
        std::memcpy(&m_vertexData[vertexByteOffset], submesh.vertices.data(), vertexSize);
        vertexByteOffset += vertexSize;
        indexByteOffset += sizeof(Model::Index) * submesh.indices.size();
    }
    CHECK(indexCounter == m_indexData.size());
}

void Raytracer::createBLAS()
{
    // src/Raytracer.cpp:1027-1157: one BLAS, one triangle geometry per submesh
    std::vector<VkrtGeometryRange> ranges;
    ranges.reserve(m_submeshIndexInfos.size());
    for (const SubmeshIndexInfo& info : m_submeshIndexInfos)
    {
        ranges.push_back(VkrtGeometryRange{info.triangleCount, info.maxVertex, info.indexByteOffset});
    }
    VkrtBuildSizes sizes{};
    VKRT_CHECK(m_device, vkrt_blas_sizes(m_device, ranges.data(), ui32Size(ranges), m_vertexData.size(), m_indexData.size(), &sizes));
    VKRT_CHECK(m_device, vkrt_blas_build(m_device, m_vertexData.data(), m_vertexData.size(), sizeof(Model::Vertex), m_indexData.data(),
                                         m_indexData.size(), ranges.data(), ui32Size(ranges), VKRT_BUILD_DEFAULT, &m_blas));
    VKRT_CHECK(m_device, vkrt_blas_info(m_device, m_blas, &m_blasInfo));
    m_vertexData.clear();
    m_vertexData.shrink_to_fit();
    m_indexData.clear();
    m_indexData.shrink_to_fit();
}

void Raytracer::createTLAS()
{
    // src/Raytracer.cpp:1159-1283: one instance, scale 0.01, mask 0xFF, cull disabled
    if (m_instances.empty())
    {
        VkrtInstance instance{};
        const float matrixData[12] = {0.01f, 0.0f, 0.0f, 0.0f, 0.0f, 0.01f, 0.0f, 0.0f, 0.0f, 0.0f, 0.01f, 0.0f};
        std::memcpy(instance.transform, matrixData, sizeof(matrixData));
        instance.instanceCustomIndex = 0;
        instance.mask = 0xFF;
        instance.instanceShaderBindingTableRecordOffset = 0;
        instance.flags = 0x1; // VK_GEOMETRY_INSTANCE_TRIANGLE_FACING_CULL_DISABLE_BIT_KHR
        m_instances.push_back(instance);
    }
    for (VkrtInstance& instance : m_instances)
    {
        instance.accelerationStructureReference = vkrt_blas_reference(m_blas);
    }
    vkrt_tlas_destroy(m_device, m_tlas);
    m_tlas = nullptr;
    VKRT_CHECK(m_device, vkrt_tlas_build(m_device, m_instances.data(), ui32Size(m_instances), &m_tlas));
}

void Raytracer::setInstances(const std::vector<VkrtInstance>& instances)
{
    m_instances = instances;
    createTLAS();
}

// >>> reference code, verbatim: src/Raytracer.cpp:1412-1427 (see PROVENANCE at the top of this file)
void Raytracer::updateMaterialIndexDescriptorSet()
{
    // src/Raytracer.cpp:1412-1427
    std::vector<SubmeshInfo> submeshInfos(m_model->submeshes.size());
    int indexBufferOffset = 0;
    for (size_t i = 0; i < m_model->submeshes.size(); ++i)
    {
        const Model::Submesh& submesh = m_model->submeshes[i];
        submeshInfos[i].baseColorTextureIndex = m_model->materials[submesh.material].baseColor;
        submeshInfos[i].normalTextureIndex = m_model->materials[submesh.material].normalImage;
        submeshInfos[i].metallicRoughnessTextureIndex = m_model->materials[submesh.material].metallicRoughnessImage;
        submeshInfos[i].indexBufferOffset = indexBufferOffset;
        indexBufferOffset += static_cast<int>(submesh.indices.size() / 3);
        // missing normal / metallic-roughness fall back to image 0 "to avoid crashes" (src/Raytracer.cpp:1424-1426)
        submeshInfos[i].normalTextureIndex = std::max(submeshInfos[i].normalTextureIndex, 0);
        submeshInfos[i].metallicRoughnessTextureIndex = std::max(submeshInfos[i].metallicRoughnessTextureIndex, 0);
    }
    VKRT_CHECK(m_device, vkrt_materials_set(m_device, reinterpret_cast<const VkrtSubmeshInfo*>(submeshInfos.data()), ui32Size(submeshInfos)));
}

void Raytracer::readback(std::vector<uint8_t>& bgra8)
{
    bgra8.resize(static_cast<size_t>(m_context.getWidth()) * m_context.getHeight() * 4);
    const std::vector<void*>& images = m_context.getSwapchainImages();
    if (!images.empty())
    {
        // the last frame was presented: take it from its ring image once the copy has landed
        m_context.waitIdle();
        memcpy(bgra8.data(), images[m_context.getLastPresentedImage()], bgra8.size());
        return;
    }
    VKRT_CHECK(m_device, vkrt_frame_read(m_device, m_frame, VKRT_AOV_BGRA8, bgra8.data(), 0));
}

void Raytracer::readbackAov(int aov, void* dst)
{
    VKRT_CHECK(m_device, vkrt_frame_read(m_device, m_frame, aov, dst, 0));
}

bool Raytracer::saveFrame(const std::string& path)
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

VkrtStats Raytracer::getStats()
{
    VkrtStats stats{};
    VKRT_CHECK(m_device, vkrt_stats_get(m_device, &stats));
    return stats;
}
