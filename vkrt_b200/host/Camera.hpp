// Camera.hpp — same public interface as the reference's Camera (src/Camera.hpp:6-33).  The only addition is the
// aspect ratio argument: the reference bakes 1600x1200 into the projection (src/Camera.cpp:10); the headless
// renderer takes the frame size at run time.
#pragma once

#include "Utils.hpp"

class Camera final
{
public:
    explicit Camera(float aspectRatio = static_cast<float>(c_windowWidth) / static_cast<float>(c_windowHeight));

    vkrt::vec3 getPosition() const;
    vkrt::vec3 getForward() const;
    vkrt::vec3 getLeft() const;
    vkrt::vec3 getUp() const;
    void setPosition(const vkrt::vec3& pos);
    void setRotation(const vkrt::vec3& rot);
    void translate(const vkrt::vec3& translation);
    void rotate(const vkrt::vec3& axis, float amount);
    void setAspectRatio(float aspectRatio);

    const vkrt::mat4& getViewMatrix() const;
    const vkrt::mat4& getProjectionMatrix() const;

private:
    void updateViewMatrix();

    vkrt::vec3 m_position = vkrt::vec3(0.0f, 0.0f, 0.0f);
    vkrt::vec3 m_rotation = vkrt::vec3(0.0f, 0.0f, 0.0f);
    vkrt::mat4 m_viewMatrix;
    vkrt::mat4 m_projectionMatrix;
};

// Fills the 256-byte uniform block exactly as Raytracer::update does (src/Raytracer.cpp:224-235).
void fillUniforms(const Camera& camera, const float lightPositions[4][4], VkrtUniforms* out);
// The reference's fixed lights (src/Raytracer.cpp:26-31).
extern const float c_lightPositions[4][4];
