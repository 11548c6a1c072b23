// DetectionIO.cpp — marker_detections.json in the reference's format
// (read: src/DetectionIO.cpp:12-58, write: :61-153; schema README.md:153-216).
#include "visual_marker_mapping/DetectionIO.h"
#include "PtreeJson.h"

#include <initializer_list>
#include <stdexcept>
#include <vector>

namespace visual_marker_mapping
{
namespace
{
void requireSections(const std::vector<std::string>& seen, std::initializer_list<const char*> keys)
{
    for (const char* key : keys) {
        bool found = false;
        for (const std::string& s : seen) found = found || s == key;
        if (!found) throw std::runtime_error(std::string("No such node (") + key + ")");   // ptree::get_child
    }
}
}

// Streamed: one image / tag / tag observation of the file is held as a tree at a time (a site-sized file has
// 4 M tag observations, > 1 GB of text).  Of a repeated top-level key every occurrence is read, not just the first.
DetectionResult readDetectionResult(const std::string& filename)
{
    DetectionResult result;
    std::vector<double> values;
    const std::vector<std::string> seen = vmm_json::read_json_sections(filename, [&](const std::string& section, const std::string&, vmm_json::Node& n) {
        if (section == "images") {
            TagImg img;
            img.imageId = n.get_int("id");
            img.filename = n.get_string("filename");
            result.images.push_back(img);
        } else if (section == "tags") {
            Tag tag;
            tag.tagId = n.get_int("id");
            tag.tagType = n.get_string("tag_type");
            tag.width = n.get_double("width");
            tag.height = n.get_double("height");
            result.tags.push_back(tag);
        } else if (section == "tag_observations") {
            TagObservation obs;
            obs.corners.reserve(4);
            for (const auto& corner : n.get_child("observations").children) {
                values.clear();
                for (const auto& value : corner.second.children) values.push_back(value.second.value_double());
                if (values.size() != 2) throw std::runtime_error("Unexpected number of values");   // DetectionIO.cpp:45-46
                obs.corners.push_back(Vector2d(values[0], values[1]));
            }
            if (obs.corners.size() != 4) throw std::runtime_error("Unexpected number of values");   // :50-51
            obs.imageId = n.get_int("image_id");
            obs.tagId = n.get_int("tag_id");
            result.tagObservations.push_back(std::move(obs));
        }
    });
    requireSections(seen, { "images", "tags", "tag_observations" });
    return result;
}

bool writeDetectionResult(const DetectionResult& result, const std::string& filename)
{
    // image paths are stored relative to the directory of the json file (DetectionIO.cpp:65-78)
    std::string base;
    const size_t slash = filename.find_last_of('/');
    if (slash != std::string::npos) base = filename.substr(0, slash + 1);
    vmm_json::Writer w(filename);
    w.begin_array("images");
    for (const auto& img : result.images) {
        std::string name = img.filename;
        if (!base.empty() && name.compare(0, base.size(), base) == 0) name = name.substr(base.size());
        w.begin_object();
        w.leaf("filename", name);
        w.leaf("id", img.imageId);
        w.end();
    }
    w.end();
    w.begin_array("tags");
    for (const auto& tag : result.tags) {
        w.begin_object();
        w.leaf("id", tag.tagId);
        w.leaf("tag_type", tag.tagType);
        w.leaf("width", tag.width);
        w.leaf("height", tag.height);
        w.end();
    }
    w.end();
    w.begin_array("tag_observations");
    for (const auto& obs : result.tagObservations) {
        w.begin_object();
        w.leaf("image_id", obs.imageId);
        w.leaf("tag_id", obs.tagId);
        w.begin_array("observations");
        for (size_t i = 0; i < 4 && i < obs.corners.size(); ++i) {
            w.begin_array();
            w.value(obs.corners[i].x());
            w.value(obs.corners[i].y());
            w.end();
        }
        w.end();
        w.end();
    }
    w.end();
    w.close();
    return true;
}
}
