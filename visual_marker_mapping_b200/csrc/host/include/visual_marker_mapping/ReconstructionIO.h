// replaces include/visual_marker_mapping/ReconstructionIO.h:13-20 (reconstruction.json, src/ReconstructionIO.cpp:57-145)
#ifndef VISUAL_MARKER_MAPPING_RECONSTRUCTIONIO_H_
#define VISUAL_MARKER_MAPPING_RECONSTRUCTIONIO_H_
#include "visual_marker_mapping/TagReconstructor.h"
#include <map>
#include <string>
namespace visual_marker_mapping
{
void exportReconstructions(const std::string& outputPath, const std::map<int, ReconstructedTag>& reconstructedTags,
    const std::map<int, Camera>& reconstructedCameras, const CameraModel& camModel);
// unlike the reference's reader (which drops the camera id, src/CameraUtilities.cpp:21-28) cameras keep their ids
void parseReconstructions(const std::string& inputPath, std::map<int, ReconstructedTag>& reconstructedTags,
    std::map<int, Camera>& reconstructedCameras, CameraModel& camModel);
}
#endif
