// Context.cpp — see Context.hpp.
#include "Context.hpp"
#include "Utils.hpp"

#include <chrono>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>

namespace
{
const uint32_t c_swapchainImageCount = 3; // src/VulkanUtils.hpp:26
}

Context::Context(const Options& options) : m_options(options)
{
    const int r = vkrt_context_create(m_options.device, &m_device);
    if (r != 0)
    {
        printf("Abort. vkrt_context_create failed: %s\n", vkrt_last_error(nullptr));
        abort();
    }
    initCommunicator();
    // "W:150,Z:150"
    std::stringstream ss(m_options.keyScript);
    std::string item;
    while (std::getline(ss, item, ','))
    {
        const size_t colon = item.find(':');
        if (item.empty() || colon == std::string::npos)
        {
            continue;
        }
        m_script.push_back(ScriptEntry{std::toupper(item[0]), std::atoi(item.c_str() + colon + 1)});
    }
}

Context::~Context()
{
    for (void* image : m_swapchainImages)
    {
        vkrt_host_free(m_device, image);
    }
    vkrt_context_destroy(m_device);
}

void Context::initCommunicator()
{
    if (m_options.worldSize <= 1)
    {
        return;
    }
    uint8_t id[VKRT_UNIQUE_ID_BYTES];
    CHECK(!m_options.commFile.empty());
    if (m_options.rank == 0)
    {
        VKRT_CHECK(nullptr, vkrt_comm_unique_id(id));
        std::ofstream tmp(m_options.commFile + ".tmp", std::ios::binary);
        tmp.write(reinterpret_cast<const char*>(id), sizeof(id));
        tmp.close();
        std::rename((m_options.commFile + ".tmp").c_str(), m_options.commFile.c_str());
    }
    else
    {
        for (int attempt = 0;; ++attempt)
        {
            std::ifstream f(m_options.commFile, std::ios::binary);
            if (f && f.read(reinterpret_cast<char*>(id), sizeof(id)))
            {
                break;
            }
            CHECK(attempt < 6000);
            std::this_thread::sleep_for(std::chrono::milliseconds(10));
        }
    }
    VKRT_CHECK(m_device, vkrt_comm_init(m_device, id, m_options.rank, m_options.worldSize));
}

void Context::scriptStep()
{
    if (m_scriptIndex >= m_script.size())
    {
        if (m_heldKey >= 0)
        {
            pushKeyEvent(m_heldKey, GLFW_RELEASE);
            m_heldKey = -1;
        }
        return;
    }
    const ScriptEntry& e = m_script[m_scriptIndex];
    if (m_scriptFrame == 0)
    {
        if (m_heldKey >= 0)
        {
            pushKeyEvent(m_heldKey, GLFW_RELEASE);
        }
        pushKeyEvent(e.key, GLFW_PRESS);
        m_heldKey = e.key;
    }
    if (++m_scriptFrame >= e.frames)
    {
        m_scriptFrame = 0;
        ++m_scriptIndex;
    }
}

bool Context::update()
{
    if (m_shouldQuit || (m_options.frames >= 0 && m_frame >= (uint64_t)m_options.frames))
    {
        return false;
    }
    scriptStep();
    ++m_frame;
    return true;
}

std::vector<Context::KeyEvent> Context::getKeyEvents()
{
    std::vector<KeyEvent> events;
    events.swap(m_keyEvents);
    for (const KeyEvent& e : events)
    {
        if (e.key == GLFW_KEY_ESCAPE && e.action == GLFW_PRESS)
        {
            m_shouldQuit = true; // src/Context.cpp:247-254
        }
    }
    return events;
}

void Context::pushKeyEvent(int key, int action)
{
    m_keyEvents.push_back(KeyEvent{key, action});
}

uint32_t Context::acquireNextSwapchainImage()
{
    if (m_imageIndex < m_pending.size() && m_pending[m_imageIndex].frame)
    {
        VKRT_CHECK(m_device, vkrt_frame_present_wait(m_device, m_pending[m_imageIndex].frame, m_pending[m_imageIndex].ticket));
        m_pending[m_imageIndex].frame = nullptr;
    }
    return m_imageIndex;
}

void Context::submitCommandBuffers(VkrtFrame* presented)
{
    if (presented)
    {
        if (m_swapchainImages.empty())
        {
            m_swapchainImages.assign(c_swapchainImageCount, nullptr);
            m_pending.assign(c_swapchainImageCount, Pending{});
            for (void*& image : m_swapchainImages)
            {
                VKRT_CHECK(m_device, vkrt_host_alloc(m_device, static_cast<size_t>(m_options.width) * m_options.height * 4, &image));
            }
        }
        Pending& p = m_pending[m_imageIndex];
        VKRT_CHECK(m_device, vkrt_frame_present(m_device, presented, m_swapchainImages[m_imageIndex], 0, &p.ticket));
        p.frame = presented;
        m_lastPresented = m_imageIndex;
    }
    m_imageIndex = (m_imageIndex + 1) % c_swapchainImageCount;
}

void Context::waitIdle()
{
    for (Pending& p : m_pending)
    {
        if (p.frame)
        {
            VKRT_CHECK(m_device, vkrt_frame_present_wait(m_device, p.frame, p.ticket));
            p.frame = nullptr;
        }
    }
}
