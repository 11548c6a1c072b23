#ifndef VMM_CAMERA_MODEL_JSON_H_
#define VMM_CAMERA_MODEL_JSON_H_
#include "PtreeJson.h"
#include "visual_marker_mapping/CameraModel.h"
namespace visual_marker_mapping
{
CameraModel cameraModelFromJson(const vmm_json::Node& n);
vmm_json::Node cameraModelToJson(const CameraModel& m);
}
#endif
