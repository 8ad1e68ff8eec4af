// Context.hpp — headless stand-in for the reference's Context (src/Context.hpp:8-69).  The window, Vulkan instance,
// swapchain and GLFW input are replaced by: a libvkrt_b200 context on one CUDA device, a frame budget that ends the
// render loop (instead of the window-close flag), and a scripted key-event queue that feeds the same
// Raytracer::updateCamera logic the reference drives from the keyboard (src/Raytracer.cpp:273-324).
#pragma once

#include "../../include/vkrt.h"

#include <cstdint>
#include <string>
#include <vector>

// GLFW constants used by the reference's input handling (values from GLFW/glfw3.h)
constexpr int GLFW_RELEASE = 0;
constexpr int GLFW_PRESS = 1;
constexpr int GLFW_REPEAT = 2;
constexpr int GLFW_KEY_A = 65, GLFW_KEY_C = 67, GLFW_KEY_D = 68, GLFW_KEY_E = 69, GLFW_KEY_Q = 81, GLFW_KEY_S = 83, GLFW_KEY_W = 87,
              GLFW_KEY_Z = 90, GLFW_KEY_ESCAPE = 256;

class Context final
{
public:
    struct KeyEvent
    {
        int key;
        int action;
    };

    struct Options
    {
        int device = 0;
        uint32_t width = 1600, height = 1200; // src/Utils.hpp:32-33 defaults
        int frames = 1;                       // update() returns false after this many frames (<0: run until ESC event)
        double fixedDeltaTime = 1.0 / 60.0;   // <=0: wall clock, as the reference
        int rank = 0, worldSize = 1;          // multi-GPU tile partition (NCCL)
        std::string commFile;                 // where rank 0 publishes the NCCL unique id
        std::string keyScript;                // e.g. "W:150,Z:150" = hold W for 150 frames, then Z for 150
    };

    explicit Context(const Options& options);
    ~Context();

    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;

    VkrtContext* getDevice() const { return m_device; }
    const Options& getOptions() const { return m_options; }
    uint32_t getWidth() const { return m_options.width; }
    uint32_t getHeight() const { return m_options.height; }
    double getFixedDeltaTime() const { return m_options.fixedDeltaTime; }

    bool update();                          // src/Context.cpp:122-127 (poll events, report whether to keep running)
    std::vector<KeyEvent> getKeyEvents();   // src/Context.cpp:129-139 (drains the queue)
    void pushKeyEvent(int key, int action);
    // src/Context.cpp:141-152: 3 images in flight -> ring index; waits until the image's previous present has landed
    uint32_t acquireNextSwapchainImage();
    // src/Context.cpp:154-180: submit + present.  Headless: the finished image of `presented` (rank 0; nullptr = nothing
    // to show) is copied asynchronously into the acquired ring image (page-locked host memory), then the ring advances.
    void submitCommandBuffers(VkrtFrame* presented = nullptr);
    const std::vector<void*>& getSwapchainImages() const { return m_swapchainImages; } // BGRA8, width*height*4 each
    uint32_t getLastPresentedImage() const { return m_lastPresented; }
    void waitIdle(); // every queued present has landed (vkDeviceWaitIdle for the presentation ring)
    uint64_t getFrameNumber() const { return m_frame; }

private:
    void initCommunicator();
    void scriptStep();

    Options m_options;
    VkrtContext* m_device = nullptr;
    std::vector<KeyEvent> m_keyEvents;
    struct ScriptEntry
    {
        int key;
        int frames;
    };
    std::vector<ScriptEntry> m_script;
    size_t m_scriptIndex = 0;
    int m_scriptFrame = 0;
    int m_heldKey = -1;
    uint64_t m_frame = 0;
    uint32_t m_imageIndex = 0;
    uint32_t m_lastPresented = 0;
    std::vector<void*> m_swapchainImages;
    struct Pending
    {
        VkrtFrame* frame = nullptr;
        uint64_t ticket = 0;
    };
    std::vector<Pending> m_pending; // per ring image: the present that last wrote it
    bool m_shouldQuit = false;
};
