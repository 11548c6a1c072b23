// visual_marker_mapping — command line tool with the reference's flag surface (src/main_mapping.cpp:16-96):
//   --project_path <dir>   reads <dir>/camera_intrinsics.json and <dir>/marker_detections.json,
//                          writes <dir>/reconstruction.json
//   --start_tag_id <id>    origin tag (default: lowest id)
// Like the reference the tool starts from the detections alone: incremental PnP initialisation with a bundle
// adjustment per added image (src/TagReconstructor.cpp:146-267; --wave: per round of images), then the closing
// sequence BA(Huber) -> cull 2 px -> BA (:271-277).  Extension: when --initial_reconstruction <json> is given its
// poses are taken as the start and only the closing sequence runs (never picked up implicitly: the same directory
// gives the same run as with the reference tool).
#include "visual_marker_mapping/CameraUtilities.h"
#include "visual_marker_mapping/DetectionIO.h"
#include "visual_marker_mapping/ReconstructionIO.h"
#include "visual_marker_mapping/TagReconstructor.h"

#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <thread>

namespace
{
struct Params
{
    std::string projectPath, initialReconstruction;
    int startTagId = -1, device = 0;
    bool overwrite = false, parity = false, help = false, wave = false;
};

Params loadParameters(int argc, char** argv)
{
    Params p;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto value = [&](const char* name) -> std::string {
            const std::string pre = std::string(name) + "=";
            if (a.compare(0, pre.size(), pre) == 0) return a.substr(pre.size());
            if (i + 1 >= argc) throw std::runtime_error(std::string("the required argument for option '") + name + "' is missing");
            return argv[++i];
        };
        if (a == "--help" || a == "-h") p.help = true;
        else if (a.compare(0, 14, "--project_path") == 0) p.projectPath = value("--project_path");
        else if (a.compare(0, 14, "--start_tag_id") == 0) p.startTagId = std::atoi(value("--start_tag_id").c_str());
        else if (a.compare(0, 24, "--initial_reconstruction") == 0) p.initialReconstruction = value("--initial_reconstruction");
        else if (a.compare(0, 8, "--device") == 0) p.device = std::atoi(value("--device").c_str());
        else if (a == "--overwrite") p.overwrite = true;
        else if (a == "--parity") p.parity = true;
        else if (a == "--wave") p.wave = true;
        else throw std::runtime_error("unrecognised option '" + a + "'");
    }
    if (!p.help && p.projectPath.empty()) throw std::runtime_error("the option '--project_path' is required but missing");
    return p;
}

bool exists(const std::string& f) { return std::ifstream(f.c_str()).good(); }
}

int main(int argc, char** argv)
{
    try {
        const Params p = loadParameters(argc, argv);
        if (p.help) {
            std::cout << "Usage: visual_marker_mapping --project_path <dir> [--start_tag_id <id>]\n"
                         "         [--initial_reconstruction <json>] [--device <cuda device>] [--overwrite] [--parity] [--wave]\n";
            return 0;
        }
        const size_t numThreads = std::thread::hardware_concurrency() ? std::thread::hardware_concurrency() : 4;
        const std::string out = p.projectPath + "/reconstruction.json";
        if (exists(out) && !p.overwrite) {   // the reference asks on stdin (src/main_mapping.cpp:57-73)
            std::cout << "Output file exists: " << out << std::endl << "Do you want to overwrite it? (y/n): ";
            char c = 0;
            std::cin >> c;
            if (c != 'y') return 0;
        }
        using namespace visual_marker_mapping;
        const CameraModel camModel = readCameraModel(p.projectPath + "/camera_intrinsics.json");
        DetectionResult detections = readDetectionResult(p.projectPath + "/marker_detections.json");
        std::map<int, ReconstructedTag> tags;
        std::map<int, Camera> cams;
        CameraModel ignored = camModel;
        const std::string init = p.initialReconstruction;
        const bool haveInit = !init.empty();
        if (haveInit) std::cout << "Initial reconstruction from " << init << ": closing sequence only" << std::endl;
        if (haveInit) parseReconstructions(init, tags, cams, ignored);
        TagReconstructor reconstructor(std::move(detections));
        reconstructor.setCameraModel(camModel);
        reconstructor.setOriginTagId(p.startTagId);
        reconstructor.setDevice(p.device);
        reconstructor.setParityMode(p.parity);
        reconstructor.setWaveMode(p.wave);
        if (haveInit) reconstructor.setInitialReconstruction(tags, cams);
        reconstructor.startReconstruction(numThreads);
        exportReconstructions(out, reconstructor.getReconstructedTags(), reconstructor.getReconstructedCameras(), reconstructor.getCameraModel());
        std::cout << "Wrote " << out << std::endl;
    } catch (const std::exception& ex) {
        std::cerr << "Error: " << ex.what() << std::endl;   // the reference also returns 0 here (src/main_mapping.cpp:90-95)
    }
    return 0;
}
