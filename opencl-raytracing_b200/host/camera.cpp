// host/camera.cpp — the camera behind include/camera.h.  Behaviour (and float rounding) of the reference's
// src/camera.cpp; structure is this repository's: a Controls record, a derived Frame, and the 12-float block built on
// demand.  Citations "ref:" point into the reference's src/camera.cpp.
#include "../../include/camera.h"

#include <cmath>

namespace {

// ref:16-22.  The two zoom limits are fields of view in degrees: zooming IN narrows the angle down to 10, OUT up to 90.
struct Tuning {
    float crawl = 0.3f, walk = 1.0f, run = 5.0f;      // units per second
    float degrees_per_pixel = 0.2f;                    // at the narrowest field of view, scaled by fov / narrowest
    float narrowest = 10.0f, widest = 90.0f, degrees_per_scroll = 0.5f;
    float pitch_limit = 89.0f;
};
const Tuning kTuning{};

// The two spellings of degrees -> radians the reference uses for the field of view.  Both multiply in double (M_PI is
// a double) and round once to float; they differ in the last bits, and each is kept where the reference has it.
float fov_radians_by_division(float degrees) { return (float)(degrees * M_PI / 180.0f); }             // ref:47 (constructor)
float fov_radians_by_reciprocal(float degrees) { return (float)(degrees * M_PI * 0.0055556f); }       // ref:40 (setFov)

}  // namespace

Camera::Camera(int camera_fov, float camera_aspect, const glm::vec3& pos, float y, float p) {
    ctl.fov_degrees = (float)camera_fov;
    ctl.aspect = camera_aspect;
    ctl.units_per_second = kTuning.crawl;              // ref:46: a fresh camera moves at the slow speed
    ctl.yaw_degrees = y;
    ctl.pitch_degrees = p;
    ctl.eye = pos;
    open(fov_radians_by_division(ctl.fov_degrees));
}

// ref:26-33: forward from yaw / pitch (yaw 0 looks down +z), right = forward x (0,-1,0), up = right x forward —
// "up" on the screen is world -y.
void Camera::orient() {
    const float pitch = glm::radians(ctl.pitch_degrees), yaw = glm::radians(ctl.yaw_degrees);
    glm::vec3 f;
    f.x = glm::cos(pitch) * glm::sin(yaw);
    f.y = glm::sin(pitch);
    f.z = glm::cos(pitch) * glm::cos(yaw);
    frame.forward = glm::normalize(f);
    frame.right = glm::normalize(glm::cross(frame.forward, glm::vec3(0.0f, -1.0f, 0.0f)));
    frame.up = glm::cross(frame.right, frame.forward);
}

// ref:41-43 / 48-50: image plane at distance 1
void Camera::open(float vertical_angle_radians) {
    frame.tan_half_vertical = std::tan(vertical_angle_radians * 0.5f);
    frame.tan_half_horizontal = ctl.aspect * frame.tan_half_vertical;
    orient();
}

// ref:53-59
void Camera::move(CameraMovementDirection dir, float dt) {
    const float distance = ctl.units_per_second * dt;
    switch (dir) {
        case FORWARD: ctl.eye += frame.forward * distance; break;
        case BACK:    ctl.eye -= frame.forward * distance; break;
        case RIGHT:   ctl.eye += frame.right * distance; break;
        case LEFT:    ctl.eye -= frame.right * distance; break;
    }
}

// ref:61-70: mouse deltas scaled with the field of view; pitch clamped short of the poles, yaw wrapped
void Camera::rotate(float x, float y) {
    ctl.yaw_degrees += x * kTuning.degrees_per_pixel * ctl.fov_degrees / kTuning.narrowest;
    ctl.pitch_degrees += y * kTuning.degrees_per_pixel * ctl.fov_degrees / kTuning.narrowest;
    if (ctl.pitch_degrees > kTuning.pitch_limit) ctl.pitch_degrees = kTuning.pitch_limit;
    else if (ctl.pitch_degrees < -kTuning.pitch_limit) ctl.pitch_degrees = -kTuning.pitch_limit;
    ctl.yaw_degrees = std::fmod(ctl.yaw_degrees, 360.0f);
    orient();
}

// ref:72-77
void Camera::zoom(float scroll) {
    float fov = ctl.fov_degrees + scroll * kTuning.degrees_per_scroll;
    if (fov < kTuning.narrowest) fov = kTuning.narrowest;
    else if (fov > kTuning.widest) fov = kTuning.widest;
    ctl.fov_degrees = fov;
    open(fov_radians_by_reciprocal(fov));
}

// ref:79-87: each switch falls back to the normal speed when released
void Camera::setFasterSpeed(bool speed_up) { ctl.units_per_second = speed_up ? kTuning.run : kTuning.walk; }
void Camera::setSlowerSpeed(bool speed_down) { ctl.units_per_second = speed_down ? kTuning.crawl : kTuning.walk; }

// ref:89-92
void Camera::setSize(float new_aspect) {
    ctl.aspect = new_aspect;
    open(fov_radians_by_reciprocal(ctl.fov_degrees));
}

// ref:34-36 + 94-118: the block the render calls read.  Corner and extents depend on the frame only, so they are
// formed here instead of being cached by every control.
float* Camera::transferData() const {
    static float block[12];
    const glm::vec3 across = frame.tan_half_horizontal * frame.right, upward = frame.tan_half_vertical * frame.up;
    const glm::vec3 corner = frame.forward - (across + upward);
    const glm::vec3 horizontal = 2.0f * frame.tan_half_horizontal * frame.right;
    const glm::vec3 vertical = 2.0f * frame.tan_half_vertical * frame.up;
    const glm::vec3* rows[4] = {&ctl.eye, &corner, &horizontal, &vertical};
    for (int r = 0; r < 4; ++r)
        for (int k = 0; k < 3; ++k) block[3 * r + k] = (*rows[r])[k];
    return block;
}
