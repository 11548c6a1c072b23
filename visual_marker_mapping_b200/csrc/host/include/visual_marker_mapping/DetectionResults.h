// DetectionResults.h — replaces include/visual_marker_mapping/DetectionResults.h:10-37.
#ifndef VISUAL_MARKER_MAPPING_DETECTIONRESULTS_H_
#define VISUAL_MARKER_MAPPING_DETECTIONRESULTS_H_
#include "visual_marker_mapping/EigenLite.h"
#include <string>
#include <vector>
namespace visual_marker_mapping
{
struct TagObservation
{
    int imageId = -1;
    int tagId = -1;
    std::vector<Vector2d> corners;   // lower-left, lower-right, upper-right, upper-left
};
struct TagImg
{
    int imageId = -1;
    std::string filename;
};
struct Tag
{
    int tagId = -1;
    std::string tagType;
    double width = 0.0;
    double height = 0.0;
};
struct DetectionResult
{
    std::vector<TagImg> images;
    std::vector<Tag> tags;
    std::vector<TagObservation> tagObservations;
};
}
#endif
