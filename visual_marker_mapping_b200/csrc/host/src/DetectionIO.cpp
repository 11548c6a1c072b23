// DetectionIO.cpp — marker_detections.json in the reference's format
// (read: src/DetectionIO.cpp:12-58, write: :61-153; schema README.md:153-216).
#include "visual_marker_mapping/DetectionIO.h"
#include "PtreeJson.h"

#include <stdexcept>

namespace visual_marker_mapping
{
DetectionResult readDetectionResult(const std::string& filename)
{
    const vmm_json::Node root = vmm_json::read_json(filename);
    DetectionResult result;
    for (const auto& it : root.get_child("images").children) {
        TagImg img;
        img.imageId = it.second.get_int("id");
        img.filename = it.second.get_string("filename");
        result.images.push_back(img);
    }
    for (const auto& it : root.get_child("tags").children) {
        Tag tag;
        tag.tagId = it.second.get_int("id");
        tag.tagType = it.second.get_string("tag_type");
        tag.width = it.second.get_double("width");
        tag.height = it.second.get_double("height");
        result.tags.push_back(tag);
    }
    for (const auto& it : root.get_child("tag_observations").children) {
        TagObservation obs;
        for (const auto& corner : it.second.get_child("observations").children) {
            std::vector<double> values;
            for (const auto& value : corner.second.children) values.push_back(value.second.value_double());
            if (values.size() != 2) throw std::runtime_error("Unexpected number of values");   // DetectionIO.cpp:45-46
            obs.corners.push_back(Vector2d(values[0], values[1]));
        }
        if (obs.corners.size() != 4) throw std::runtime_error("Unexpected number of values");   // :50-51
        obs.imageId = it.second.get_int("image_id");
        obs.tagId = it.second.get_int("tag_id");
        result.tagObservations.push_back(obs);
    }
    return result;
}

bool writeDetectionResult(const DetectionResult& result, const std::string& filename)
{
    using vmm_json::Node;
    // image paths are stored relative to the directory of the json file (DetectionIO.cpp:65-78)
    std::string base;
    const size_t slash = filename.find_last_of('/');
    if (slash != std::string::npos) base = filename.substr(0, slash + 1);
    Node root, imgs, tags, observations;
    for (const auto& img : result.images) {
        Node n;
        std::string name = img.filename;
        if (!base.empty() && name.compare(0, base.size(), base) == 0) name = name.substr(base.size());
        n.put("filename", name);
        n.put("id", img.imageId);
        imgs.push_back(n);
    }
    root.add_child("images", imgs);
    for (const auto& tag : result.tags) {
        Node n;
        n.put("id", tag.tagId);
        n.put("tag_type", tag.tagType);
        n.put("width", tag.width);
        n.put("height", tag.height);
        tags.push_back(n);
    }
    root.add_child("tags", tags);
    for (const auto& obs : result.tagObservations) {
        Node n, corners;
        n.put("image_id", obs.imageId);
        n.put("tag_id", obs.tagId);
        for (size_t i = 0; i < 4 && i < obs.corners.size(); ++i) {
            Node coord;
            coord.push_back(Node::leaf(obs.corners[i].x()));
            coord.push_back(Node::leaf(obs.corners[i].y()));
            corners.push_back(coord);
        }
        n.add_child("observations", corners);
        observations.push_back(n);
    }
    root.add_child("tag_observations", observations);
    vmm_json::write_json(filename, root);
    return true;
}
}
