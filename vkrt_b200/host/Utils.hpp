// Utils.hpp — host-side constants and abort macros; mirrors the reference's src/Utils.hpp:8-43 and the
// VK_CHECK convention of src/VulkanUtils.hpp:28-37 (print + abort), here wrapping the C ABI's error codes.
#pragma once

#include "Math.hpp"
#include "../../include/vkrt.h"

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>

#define CHECK(f)                                                           \
    do                                                                     \
    {                                                                      \
        if (!(f))                                                          \
        {                                                                  \
            printf("Abort. %s failed at %s:%d\n", #f, __FILE__, __LINE__); \
            abort();                                                       \
        }                                                                  \
    } while (false)

#define LOGE(f)                                                \
    do                                                         \
    {                                                          \
        printf("ERROR: %s at %s:%d\n", f, __FILE__, __LINE__); \
        abort();                                               \
    } while (false)

#define LOGW(f)                                                  \
    do                                                           \
    {                                                            \
        printf("WARNING: %s at %s:%d\n", f, __FILE__, __LINE__); \
    } while (false)

// The reference aborts on any failed Vulkan call; the C ABI never aborts, so the host wrapper does.
#define VKRT_CHECK(ctx, f)                                                                               \
    do                                                                                                   \
    {                                                                                                    \
        const int result__ = (f);                                                                        \
        if (result__ != 0)                                                                               \
        {                                                                                                \
            printf("Abort. %s failed at %s:%d. Result = %d (%s)\n", #f, __FILE__, __LINE__, result__,     \
                   vkrt_last_error(ctx));                                                                \
            abort();                                                                                     \
        }                                                                                                \
    } while (false)

// Window size of the reference (src/Utils.hpp:32-33); the headless port makes the frame size a runtime value and
// these are only the defaults.
const int c_windowWidth = 1600;
const int c_windowHeight = 1200;

const vkrt::vec3 c_forward(0.0f, 0.0f, -1.0f);
const vkrt::vec4 c_forwardZero(c_forward.x, c_forward.y, c_forward.z, 0.0f);
const vkrt::vec3 c_backward(0.0f, 0.0f, 1.0f);
const vkrt::vec3 c_up(0.0f, 1.0f, 0.0f);
const vkrt::vec4 c_upZero(c_up.x, c_up.y, c_up.z, 0.0f);
const vkrt::vec3 c_down(0.0f, -1.0f, 0.0f);
const vkrt::vec3 c_left(-1.0f, 0.0f, 0.0f);
const vkrt::vec4 c_leftZero(c_left.x, c_left.y, c_left.z, 0.0f);
const vkrt::vec3 c_right(1.0f, 0.0f, 0.0f);

inline vkrt::vec4 toVec4(vkrt::vec3 v, float w)
{
    return vkrt::vec4(v.x, v.y, v.z, w);
}

template<typename T>
uint32_t ui32Size(const T& container)
{
    return static_cast<uint32_t>(container.size());
}
