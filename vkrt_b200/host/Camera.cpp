// Camera.cpp — restates the reference's src/Camera.cpp:6-78 (pose -> view / projection matrices).
#include "Camera.hpp"

#include <cstring>

using namespace vkrt;

const float c_lightPositions[4][4] = {
    {6.0f, 6.0f, 0.0f, 0.0f},
    {2.0f, 5.0f, 0.0f, 0.0f},
    {-2.0f, 4.0f, 0.0f, 0.0f},
    {-6.0f, 3.0f, 0.0f, 0.0f},
};

Camera::Camera(float aspectRatio)
{
    updateViewMatrix();
    setAspectRatio(aspectRatio);
}

void Camera::setAspectRatio(float aspectRatio)
{
    // The reference passes 45.0f to glm::perspective, which takes radians (src/Camera.cpp:9,13); kept as is.
    const float fov = 45.0f;
    const float nearClipDistance = 0.1f;
    const float farClipDistance = 100.0f;
    m_projectionMatrix = perspective(fov, aspectRatio, nearClipDistance, farClipDistance);
    m_projectionMatrix[1][1] *= -1; // Vulkan clip-space Y flip (src/Camera.cpp:14)
}

vec3 Camera::getPosition() const
{
    return m_position;
}

vec3 Camera::getForward() const
{
    const vec4 v = yawPitchRoll(m_rotation.y, m_rotation.x, m_rotation.z) * c_forwardZero;
    return vec3(v.x, v.y, v.z);
}

vec3 Camera::getLeft() const
{
    const vec4 v = yawPitchRoll(m_rotation.y, m_rotation.x, m_rotation.z) * c_leftZero;
    return vec3(v.x, v.y, v.z);
}

vec3 Camera::getUp() const
{
    const vec4 v = yawPitchRoll(m_rotation.y, m_rotation.x, m_rotation.z) * c_upZero;
    return vec3(v.x, v.y, v.z);
}

void Camera::setPosition(const vec3& pos)
{
    m_position = pos;
    updateViewMatrix();
}

void Camera::setRotation(const vec3& rot)
{
    m_rotation = rot;
    updateViewMatrix();
}

void Camera::translate(const vec3& translation)
{
    m_position += translation;
    updateViewMatrix();
}

void Camera::rotate(const vec3& axis, float amount)
{
    m_rotation += axis * amount;
    updateViewMatrix();
}

const mat4& Camera::getViewMatrix() const
{
    return m_viewMatrix;
}

const mat4& Camera::getProjectionMatrix() const
{
    return m_projectionMatrix;
}

void Camera::updateViewMatrix()
{
    const vec3 forward = getForward();
    m_viewMatrix = lookAt(m_position, m_position + forward, c_up);
}

void fillUniforms(const Camera& camera, const float lightPositions[4][4], VkrtUniforms* out)
{
    const vec4 forward = toVec4(camera.getForward(), 0.0f);
    const vec4 right = toVec4(-camera.getLeft(), 0.0f);
    const vec4 up = toVec4(camera.getUp(), 0.0f);
    const vec4 position = toVec4(camera.getPosition(), 1.0f);
    const mat4 projInverse = inverse(camera.getProjectionMatrix());
    const mat4 viewInverse = inverse(camera.getViewMatrix());
    std::memcpy(out->viewInverse, &viewInverse, 64);
    std::memcpy(out->projInverse, &projInverse, 64);
    std::memcpy(out->position, &position, 16);
    std::memcpy(out->right, &right, 16);
    std::memcpy(out->up, &up, 16);
    std::memcpy(out->forward, &forward, 16);
    std::memcpy(out->lightPositions, lightPositions, 64);
}
