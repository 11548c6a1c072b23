// CameraModel.h — replaces include/visual_marker_mapping/CameraModel.h:12-34 (OpenCV pinhole +
// radial-tangential model; x right, y down, z forward).
#ifndef VISUAL_MARKER_MAPPING_CAMERAMODEL_H_
#define VISUAL_MARKER_MAPPING_CAMERAMODEL_H_
#include "visual_marker_mapping/EigenLite.h"
namespace visual_marker_mapping
{
struct CameraModel
{
    double fx = 0, fy = 0, cx = 0, cy = 0;
    double distortionCoefficients[5] = { 0, 0, 0, 0, 0 };   // k1, k2, p1, p2, k3
    int verticalResolution = 0;
    int horizontalResolution = 0;
    // replaces CameraModel::projectPoint (src/CameraModel.cpp:6-26), including its use of the already
    // distorted x inside the y formula.
    Vector2d projectPoint(const Vector3d& point3D) const;
};
}
#endif
