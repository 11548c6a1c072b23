// replaces include/visual_marker_mapping/DetectionIO.h:9-10 (marker_detections.json, src/DetectionIO.cpp:12-153)
#ifndef VISUAL_MARKER_MAPPING_DETECTIONIO_H_
#define VISUAL_MARKER_MAPPING_DETECTIONIO_H_
#include "visual_marker_mapping/DetectionResults.h"
#include <string>
namespace visual_marker_mapping
{
DetectionResult readDetectionResult(const std::string& filename);
bool writeDetectionResult(const DetectionResult& result, const std::string& filename);
}
#endif
