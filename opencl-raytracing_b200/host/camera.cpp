// host/camera.cpp — Camera of the reference (src/camera.cpp) restated; constants of src/camera.cpp:16-23.
#include "../../include/camera.h"

#include <cmath>

namespace {
const float kSpeedSlow = 0.3f, kSpeedNormal = 1.0f, kSpeedFast = 5.0f;   // src/camera.cpp:16-18
const float kMouseSensitivity = 0.2f;                                    // :19
const float kZoomMin = 90.0f, kZoomMax = 10.0f, kZoomSpeed = 0.5f;       // :20-22 (fov limits, sic)
const glm::vec3 kUp(0.0f, -1.0f, 0.0f);                                  // :23, "up" on screen is world -y
}  // namespace

// src/camera.cpp:26-37
void Camera::updateVectors() {
    const float rp = glm::radians(pitch), ry = glm::radians(yaw);
    w.x = glm::cos(rp) * glm::sin(ry);
    w.y = glm::sin(rp);
    w.z = glm::cos(rp) * glm::cos(ry);
    w = glm::normalize(w);
    u = glm::normalize(glm::cross(w, kUp));
    v = glm::cross(u, w);
    lower_left_corner = w - (half_width * u + half_height * v);
    horizontal = 2.0f * half_width * u;
    vertical = 2.0f * half_height * v;
}

// src/camera.cpp:39-44 — note the 0.0055556f (not pi/180 as in the constructor); both are kept
void Camera::setFov() {
    const float angle = (float)(fov * M_PI * 0.0055556f);
    half_height = std::tan(angle * 0.5f);
    half_width = aspect * half_height;
    updateVectors();
}

// src/camera.cpp:46-51
Camera::Camera(int camera_fov, float camera_aspect, const glm::vec3& pos, float y, float p)
    : fov((float)camera_fov), aspect(camera_aspect), speed(kSpeedSlow), yaw(y), pitch(p), position(pos) {
    const float angle = (float)(fov * M_PI / 180.0f);
    half_height = std::tan(angle * 0.5f);
    half_width = aspect * half_height;
    updateVectors();
}

// src/camera.cpp:53-59
void Camera::move(CameraMovementDirection dir, float dt) {
    const float ds = speed * dt;
    if (dir == FORWARD) position += w * ds;
    else if (dir == BACK) position -= w * ds;
    else if (dir == LEFT) position -= u * ds;
    else if (dir == RIGHT) position += u * ds;
}

// src/camera.cpp:61-70
void Camera::rotate(float x, float y) {
    yaw += x * kMouseSensitivity * fov / kZoomMax;
    pitch += y * kMouseSensitivity * fov / kZoomMax;
    if (pitch > 89.0f) pitch = 89.0f;
    else if (pitch < -89.0f) pitch = -89.0f;
    yaw = std::fmod(yaw, 360.0f);
    updateVectors();
}

// src/camera.cpp:72-77
void Camera::zoom(float scroll) {
    fov += scroll * kZoomSpeed;
    if (fov < kZoomMax) fov = kZoomMax;
    else if (fov > kZoomMin) fov = kZoomMin;
    setFov();
}

void Camera::setFasterSpeed(bool speed_up) { speed = speed_up ? kSpeedFast : kSpeedNormal; }      // :79-82
void Camera::setSlowerSpeed(bool speed_down) { speed = speed_down ? kSpeedSlow : kSpeedNormal; }  // :84-87
void Camera::setSize(float new_aspect) { aspect = new_aspect; setFov(); }                         // :89-92

// src/camera.cpp:94-118
float* Camera::transferData() const {
    static float data[12];
    const glm::vec3* src[4] = {&position, &lower_left_corner, &horizontal, &vertical};
    for (int i = 0; i < 4; ++i) { data[3 * i] = src[i]->x; data[3 * i + 1] = src[i]->y; data[3 * i + 2] = src[i]->z; }
    return data;
}
