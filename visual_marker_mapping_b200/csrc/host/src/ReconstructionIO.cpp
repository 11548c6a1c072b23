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
template <int N> vmm_json::Node vectorNode(const VectorNd<N>& v)
{
    vmm_json::Node n;
    for (int i = 0; i < N; ++i) n.push_back(vmm_json::Node::leaf(v(i)));
    return n;
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
    using vmm_json::Node;
    Node root, tags, corners, cams;
    for (const auto& it : reconstructedTags) {
        const ReconstructedTag& tag = it.second;
        Node n;
        n.put("id", tag.id);
        n.put("type", tag.tagType);
        n.put("width", tag.tagWidth);
        n.put("height", tag.tagHeight);
        n.add_child("rotation", vectorNode(tag.q));
        n.add_child("translation", vectorNode(tag.t));
        tags.push_back(n);
    }
    root.add_child("reconstructed_tags", tags);
    for (const auto& it : reconstructedTags) {
        const std::vector<Vector3d> c = it.second.computeMarkerCorners3D();
        for (int i = 0; i < 4; ++i) {
            Node n;
            n.put("marker_id", it.second.id);
            n.put("corner_index", i);
            n.add_child("coords", vectorNode(c[i]));
            corners.push_back(n);
        }
    }
    root.add_child("reconstructed_marker_corners", corners);
    for (const auto& it : reconstructedCameras) {
        Node n;
        n.put("id", it.second.cameraId);
        n.add_child("rotation", vectorNode(it.second.q));
        n.add_child("translation", vectorNode(it.second.t));
        cams.push_back(n);
    }
    root.add_child("reconstructed_cameras", cams);
    root.add_child("camera_model", cameraModelToJson(camModel));
    vmm_json::write_json(outputPath, root);
}

void parseReconstructions(const std::string& inputPath, std::map<int, ReconstructedTag>& reconstructedTags,
    std::map<int, Camera>& reconstructedCameras, CameraModel& camModel)
{
    const vmm_json::Node root = vmm_json::read_json(inputPath);
    reconstructedTags.clear();
    reconstructedCameras.clear();
    for (const auto& it : root.get_child("reconstructed_tags").children) {
        const vmm_json::Node& n = it.second;
        ReconstructedTag tag;
        tag.id = n.get_int("id");
        tag.q = nodeVector<4>(n.get_child("rotation"));
        tag.t = nodeVector<3>(n.get_child("translation"));
        tag.tagWidth = n.get_double("width");
        tag.tagHeight = n.get_double("height");
        tag.tagType = n.get_string("type");
        reconstructedTags[tag.id] = tag;
    }
    for (const auto& it : root.get_child("reconstructed_cameras").children) {
        Camera cam;
        cam.cameraId = it.second.get_int("id");
        cam.q = nodeVector<4>(it.second.get_child("rotation"));
        cam.t = nodeVector<3>(it.second.get_child("translation"));
        reconstructedCameras[cam.cameraId] = cam;
    }
    if (const vmm_json::Node* m = root.find("camera_model")) camModel = cameraModelFromJson(*m);
}
}
