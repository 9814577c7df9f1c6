// include/camera.h — the reference's Camera (include/camera.h:13-53, src/camera.cpp) with the same public
// interface.  transferData() returns the 12 floats [position, lower_left_corner, horizontal, vertical] the render
// calls take (src/camera.cpp:94-118); like the reference it points at function-static storage (not re-entrant).
#ifndef camera_h
#define camera_h

#include "rt_host_types.h"

enum CameraMovementDirection { FORWARD, BACK, LEFT, RIGHT };

class Camera {
private:
    float fov, aspect;
    float speed;
    float half_height;
    float half_width;
    float yaw, pitch;
    glm::vec3 u, v, w;
    glm::vec3 position;
    glm::vec3 horizontal, vertical, lower_left_corner;

    void updateVectors();
    void setFov();

public:
    Camera(int camera_fov, float camera_aspect, const glm::vec3& pos = glm::vec3(0.0f, 0.0f, 0.0f), float y = 0.0f, float p = 0.0f);

    void move(CameraMovementDirection dir, float dt);
    void rotate(float x, float y);
    void zoom(float scroll);

    void setFasterSpeed(bool speed_up);
    void setSlowerSpeed(bool speed_down);
    void setSize(float new_aspect);

    float* transferData() const;
};

#endif /* camera_h */
