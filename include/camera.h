// include/camera.h — pinhole camera with the PUBLIC interface of the reference's Camera (reference
// include/camera.h:20-53, src/camera.cpp), so that main.cpp-style callers compile unchanged:
//
//     Camera cam(60, 16.0f / 9.0f);  cam.move(FORWARD, dt);  cam.rotate(dx, dy);  cam.zoom(s);
//     renderer.render(&cam);                       // reads cam.transferData()
//
// Internally the camera keeps two small records — the controls' state and the orthonormal view frame derived from
// it — and builds the 12 floats the render calls take on demand.  The float arithmetic of every control is the
// reference's, operation for operation (tests/test_host_classes.py compares bit patterns with a restatement of
// src/camera.cpp on random control sequences).
#ifndef camera_h
#define camera_h

#include "rt_host_types.h"

enum CameraMovementDirection { FORWARD, BACK, LEFT, RIGHT };

class Camera {
public:
    Camera(int camera_fov, float camera_aspect, const glm::vec3& pos = glm::vec3(0.0f, 0.0f, 0.0f), float y = 0.0f, float p = 0.0f);

    // controls (src/camera.cpp:53-92)
    void move(CameraMovementDirection dir, float dt);
    void rotate(float x, float y);
    void zoom(float scroll);
    void setFasterSpeed(bool speed_up);
    void setSlowerSpeed(bool speed_down);
    void setSize(float new_aspect);

    // [eye, lower-left corner of the image plane, horizontal extent, vertical extent] (src/camera.cpp:94-118).
    // Like the reference's, the pointer refers to function-static storage: valid until the next call, not re-entrant.
    float* transferData() const;

private:
    struct Controls {            // what the user steers
        float fov_degrees;
        float aspect;
        float units_per_second;
        float yaw_degrees, pitch_degrees;
        glm::vec3 eye;
    };
    struct Frame {               // derived: view direction, screen-right, screen-up, half extents of the image plane at distance 1
        glm::vec3 forward, right, up;
        float tan_half_vertical, tan_half_horizontal;
    };
    Controls ctl;
    Frame frame;

    void orient();                               // yaw / pitch -> frame axes
    void open(float vertical_angle_radians);     // field of view -> half extents, then orient()
};

#endif /* camera_h */
