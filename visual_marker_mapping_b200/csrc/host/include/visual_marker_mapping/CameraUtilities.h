// replaces include/visual_marker_mapping/CameraUtilities.h:20 (camera_intrinsics.json, src/CameraUtilities.cpp:45-67)
#ifndef VISUAL_MARKER_MAPPING_CAMERAUTILITIES_H_
#define VISUAL_MARKER_MAPPING_CAMERAUTILITIES_H_
#include "visual_marker_mapping/CameraModel.h"
#include <string>
namespace visual_marker_mapping
{
CameraModel readCameraModel(const std::string& filename);
}
#endif
