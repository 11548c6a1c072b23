// Camera.h — replaces include/visual_marker_mapping/Camera.h:9-18 (pose maps world -> camera).
#ifndef VISUAL_MARKER_MAPPING_CAMERA_H
#define VISUAL_MARKER_MAPPING_CAMERA_H
#include "visual_marker_mapping/EigenLite.h"
namespace visual_marker_mapping
{
struct Camera
{
    int cameraId = -1;
    Vector4d q = Vector4d::UnitX();   // (w,x,y,z)
    Vector3d t = Vector3d::Zero();
};
}
#endif
