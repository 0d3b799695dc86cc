// C wrapper around the reference's own Camera class (include/camera.h, src/camera.cpp),
// compiled by ref_build.sh together with the UNMODIFIED /root/reference/src/camera.cpp.
// Test infrastructure (oracle/_ref/libref_camera.so).
#include "camera.h"
extern "C" {
void* refcam_create(float x, float y, float z) { return new Camera(glm::vec3(x, y, z)); }
void refcam_destroy(void* c) { delete (Camera*)c; }
void refcam_move(void* c, int dir, float d) { ((Camera*)c)->move((CameraMovementDirection)dir, d); }
void refcam_mouse(void* c, float x, float y) { ((Camera*)c)->processMouseMovement(x, y); }
void refcam_look(void* c, int dir) { ((Camera*)c)->look((CameraMovementDirection)dir); }
void refcam_get(void* c, float* o) {  // position, right, up, front
  Camera* cam = (Camera*)c;
  glm::vec3 v[4] = {cam->getPosition(), cam->getRightVector(), cam->getUpVector(), cam->getFrontVector()};
  for (int i = 0; i < 4; i++) { o[3 * i] = v[i].x; o[3 * i + 1] = v[i].y; o[3 * i + 2] = v[i].z; }
}
void refcam_view(void* c, int without_translation, float* o16) {
  Camera* cam = (Camera*)c;
  glm::mat4 m = without_translation ? cam->getViewingMatrixWithoutTranslation() : cam->getViewingMatrix();
  for (int i = 0; i < 16; i++) o16[i] = (&m.m[0][0])[i];
}
}
