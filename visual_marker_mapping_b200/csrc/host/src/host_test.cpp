// vmm_host_test — test driver of the host facade (tests/test_host_facade.py).
//   vmm_host_test roundtrip <project_dir> <out_dir>   re-writes the three project files through the C++ reader/writer (no GPU)
//   vmm_host_test stats <project_dir>                 GPU: reprojection statistics at the initial poses, as JSON on stdout
#include "PtreeJson.h"
#include "visual_marker_mapping/CameraUtilities.h"
#include "visual_marker_mapping/DetectionIO.h"
#include "visual_marker_mapping/ReconstructionIO.h"
#include "visual_marker_mapping/TagReconstructor.h"

#include <cstdio>
#include <iostream>
#include <string>

using namespace visual_marker_mapping;

int main(int argc, char** argv)
{
    try {
        if (argc < 3) { std::cerr << "usage: vmm_host_test roundtrip|stats ..." << std::endl; return 2; }
        const std::string mode = argv[1], dir = argv[2];
        if (mode == "roundtrip") {
            if (argc < 4) return 2;
            const std::string out = argv[3];
            const DetectionResult det = readDetectionResult(dir + "/marker_detections.json");
            writeDetectionResult(det, out + "/marker_detections.json");
            std::map<int, ReconstructedTag> tags;
            std::map<int, Camera> cams;
            CameraModel m = readCameraModel(dir + "/camera_intrinsics.json");
            parseReconstructions(dir + "/initial_reconstruction.json", tags, cams, m);
            exportReconstructions(out + "/initial_reconstruction.json", tags, cams, m);
            std::printf("images %zu tags %zu observations %zu rec_tags %zu rec_cams %zu fx %.17g\n", det.images.size(), det.tags.size(),
                det.tagObservations.size(), tags.size(), cams.size(), m.fx);
            return 0;
        }
        if (mode == "stats") {
            DetectionResult det = readDetectionResult(dir + "/marker_detections.json");
            std::map<int, ReconstructedTag> tags;
            std::map<int, Camera> cams;
            CameraModel m = readCameraModel(dir + "/camera_intrinsics.json");
            parseReconstructions(dir + "/initial_reconstruction.json", tags, cams, m);
            TagReconstructor rec(std::move(det));
            rec.setCameraModel(m);
            rec.setQuiet(true);
            rec.setInitialReconstruction(tags, cams);
            double avg = 0;
            const auto perTag = rec.computeReprojectionErrorPerTag(avg);
            const auto perImg = rec.computeReprojectionErrorPerImg();
            vmm_json::Node root, t, c;
            root.put("avg", avg);
            for (const auto& e : perTag) t.put(std::to_string(e.first), e.second);
            for (const auto& e : perImg) c.put(std::to_string(e.first), e.second);
            root.add_child("per_tag", t);
            root.add_child("per_img", c);
            std::cout << vmm_json::to_json(root);
            return 0;
        }
        return 2;
    } catch (const std::exception& ex) {
        std::cerr << "Error: " << ex.what() << std::endl;
        return 1;
    }
}
