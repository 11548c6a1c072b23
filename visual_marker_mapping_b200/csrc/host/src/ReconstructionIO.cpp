// ReconstructionIO.cpp — reconstruction.json (write: src/ReconstructionIO.cpp:12-97, read: :99-145;
// README.md:220-271).  Key order as the reference writes it: reconstructed_tags, reconstructed_marker_corners,
// reconstructed_cameras, camera_model.
#include "visual_marker_mapping/ReconstructionIO.h"
#include "CameraModelJson.h"

#include <stdexcept>

namespace visual_marker_mapping
{
namespace
{
template <int N> void writeVector(vmm_json::Writer& w, const char* key, const VectorNd<N>& v)
{
    w.begin_array(key);
    for (int i = 0; i < N; ++i) w.value(v(i));
    w.end();
}
template <int N> VectorNd<N> nodeVector(const vmm_json::Node& n)
{
    VectorNd<N> v;
    int i = 0;
    for (const auto& c : n.children) {
        if (i == N) throw std::runtime_error("Too many parameters in vector. Expected " + std::to_string(N) + "!");
        v(i++) = c.second.value_double();
    }
    if (i < N) throw std::runtime_error("Not enough many parameters in vector. Expected " + std::to_string(N) + ", got " + std::to_string(i) + "!");
    return v;
}
}

void exportReconstructions(const std::string& outputPath, const std::map<int, ReconstructedTag>& reconstructedTags,
    const std::map<int, Camera>& reconstructedCameras, const CameraModel& camModel)
{
    vmm_json::Writer w(outputPath);   // streamed: 120 k poses + 80 k corners at site size
    w.begin_array("reconstructed_tags");
    for (const auto& it : reconstructedTags) {
        const ReconstructedTag& tag = it.second;
        w.begin_object();
        w.leaf("id", tag.id);
        w.leaf("type", tag.tagType);
        w.leaf("width", tag.tagWidth);
        w.leaf("height", tag.tagHeight);
        writeVector(w, "rotation", tag.q);
        writeVector(w, "translation", tag.t);
        w.end();
    }
    w.end();
    w.begin_array("reconstructed_marker_corners");
    for (const auto& it : reconstructedTags) {
        const std::vector<Vector3d> c = it.second.computeMarkerCorners3D();
        for (int i = 0; i < 4; ++i) {
            w.begin_object();
            w.leaf("marker_id", it.second.id);
            w.leaf("corner_index", i);
            writeVector(w, "coords", c[i]);
            w.end();
        }
    }
    w.end();
    w.begin_array("reconstructed_cameras");
    for (const auto& it : reconstructedCameras) {
        w.begin_object();
        w.leaf("id", it.second.cameraId);
        writeVector(w, "rotation", it.second.q);
        writeVector(w, "translation", it.second.t);
        w.end();
    }
    w.end();
    w.node("camera_model", cameraModelToJson(camModel));
    w.close();
}

void parseReconstructions(const std::string& inputPath, std::map<int, ReconstructedTag>& reconstructedTags,
    std::map<int, Camera>& reconstructedCameras, CameraModel& camModel)
{
    reconstructedTags.clear();
    reconstructedCameras.clear();
    vmm_json::Node model;
    const std::vector<std::string> seen = vmm_json::read_json_sections(inputPath, [&](const std::string& section, const std::string& key, vmm_json::Node& n) {
        if (section == "reconstructed_tags") {
            ReconstructedTag tag;
            tag.id = n.get_int("id");
            tag.q = nodeVector<4>(n.get_child("rotation"));
            tag.t = nodeVector<3>(n.get_child("translation"));
            tag.tagWidth = n.get_double("width");
            tag.tagHeight = n.get_double("height");
            tag.tagType = n.get_string("type");
            reconstructedTags[tag.id] = tag;
        } else if (section == "reconstructed_cameras") {
            Camera cam;
            cam.cameraId = n.get_int("id");
            cam.q = nodeVector<4>(n.get_child("rotation"));
            cam.t = nodeVector<3>(n.get_child("translation"));
            reconstructedCameras.emplace(cam.cameraId, cam);   // a repeated id keeps the first entry, as the reference's emplace does (src/ReconstructionIO.cpp:106); tags: last wins (:119-128)
        } else if (section == "camera_model") {
            model.add_child(key, std::move(n));
        }
    });
    for (const char* key : { "reconstructed_tags", "reconstructed_cameras", "camera_model" }) {   // get_child throws on each (src/ReconstructionIO.cpp:141-143)
        bool found = false;
        for (const std::string& s : seen) found = found || s == key;
        if (!found) throw std::runtime_error(std::string("No such node (") + key + ")");
    }
    for (const std::string& s : seen)
        if (s == "camera_model") { camModel = cameraModelFromJson(model); break; }
}
}
