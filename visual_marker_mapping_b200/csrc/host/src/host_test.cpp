// vmm_host_test — test driver of the host facade (tests/test_host_facade.py).
//   vmm_host_test roundtrip <project_dir> <out_dir>   re-writes the three project files through the C++ reader/writer (no GPU)
//   vmm_host_test stats <project_dir>                 GPU: reprojection statistics at the initial poses, as JSON on stdout
//   vmm_host_test bigio <out_dir> <n>                  streamed write + read-back of n synthetic tag observations; prints the peak RSS (no GPU)
//   vmm_host_test pnp <cases.txt>                     PnP.h on text cases (no GPU): "case <n> <robust> fx fy cx cy k1 k2 p1 p2 k3" then
//                                                     n lines "X Y Z u v"; prints "R(9 row-major) t(3) rms inliers" per case
#include "PtreeJson.h"
#include "visual_marker_mapping/CameraUtilities.h"
#include "visual_marker_mapping/DetectionIO.h"
#include "visual_marker_mapping/ReconstructionIO.h"
#include "visual_marker_mapping/PnP.h"
#include "visual_marker_mapping/TagReconstructor.h"

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <string>
#include <sys/resource.h>
#include <vector>

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
        if (mode == "flatten") {   // vmm_host_test flatten <n_cam> <n_tag> <n_obs>: TagReconstructor::flatProblem against a plain restatement (no GPU)
            if (argc < 5) return 2;
            const int nCam = std::atoi(argv[2]), nTag = std::atoi(argv[3]);
            const size_t nObs = (size_t)std::atoll(argv[4]);
            DetectionResult det;
            std::map<int, ReconstructedTag> tags;
            std::map<int, Camera> cams;
            // sparse, unordered ids; every 7th image and every 11th tag are not reconstructed (their observations are skipped)
            for (int i = 0; i < nCam; ++i) {
                const int id = 3 * i + 5;
                det.images.push_back(TagImg{ id, "img" });
                if (i % 7 == 3) continue;
                Camera c;
                c.cameraId = id;
                c.t = Vector3d(i, 0.5 * i, 1.0);
                cams[id] = c;
            }
            for (int j = 0; j < nTag; ++j) {
                const int id = (j % 2 ? 1000000 : 0) + 2 * j;
                det.tags.push_back(Tag{ id, "apriltag_36h11", 0.1285, 0.1295 });
                if (j % 11 == 4) continue;
                ReconstructedTag t;
                t.id = id;
                t.tagType = "apriltag_36h11";
                t.tagWidth = 0.1 + 1e-4 * j;
                t.tagHeight = 0.2;
                t.t = Vector3d(0.0, j, -1.0);
                tags[id] = t;
            }
            det.tagObservations.resize(nObs);
            uint64_t state = 88172645463325252ull;
            const auto rnd = [&state]() { state ^= state << 13; state ^= state >> 7; state ^= state << 17; return state; };
            for (size_t k = 0; k < nObs; ++k) {
                TagObservation& o = det.tagObservations[k];
                o.imageId = 3 * (int)(rnd() % (uint64_t)nCam) + 5;
                const int j = (int)(rnd() % (uint64_t)nTag);
                o.tagId = (j % 2 ? 1000000 : 0) + 2 * j;
                for (int c = 0; c < 4; ++c) o.corners.push_back(Vector2d((double)(rnd() % 6000), (double)(rnd() % 4000)));
            }
            const std::vector<TagObservation> obs = det.tagObservations;
            TagReconstructor rec(std::move(det));
            rec.setQuiet(true);
            rec.setInitialReconstruction(tags, cams);
            rec.setOriginTagId(tags.begin()->first);
            const auto t0 = std::chrono::steady_clock::now();
            rec.flatProblem();
            const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            const auto t2 = std::chrono::steady_clock::now();
            const TagReconstructor::FlatProblem& f = rec.flatProblem();   // second call: buffers reused
            const double ms2 = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t2).count();
            // plain restatement: std::map lookups, observation by observation
            const auto t1 = std::chrono::steady_clock::now();
            std::map<int, int> camIndex, tagIndex;
            for (const auto& c : cams) { const int n = (int)camIndex.size(); camIndex[c.first] = n; }
            for (const auto& t : tags) { const int n = (int)tagIndex.size(); tagIndex[t.first] = n; }
            size_t bad = (f.camIds.size() != cams.size()) + (f.tagIds.size() != tags.size()) + (f.fixedTag != 0), n = 0;
            for (size_t k = 0; k < obs.size(); ++k) {
                const auto c = camIndex.find(obs[k].imageId);
                const auto t = tagIndex.find(obs[k].tagId);
                if (c == camIndex.end() || t == tagIndex.end()) continue;
                if (n < f.toCam.size()) {
                    bad += f.toCam[n] != c->second || f.toTag[n] != t->second || f.toSource[n] != k;
                    for (int i = 0; i < 4; ++i) bad += f.toUV[8 * n + 2 * i] != obs[k].corners[i].x() || f.toUV[8 * n + 2 * i + 1] != obs[k].corners[i].y();
                }
                ++n;
            }
            const double ms_ref = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t1).count();
            bad += n != f.toCam.size() || f.toTag.size() != n || f.toUV.size() != 8 * n || f.toSource.size() != n;
            size_t j = 0;
            for (const auto& t : tags) { bad += f.tagIds[j] != t.first || f.tagWH[2 * j] != t.second.tagWidth || f.tagT[3 * j + 1] != t.second.t(1); ++j; }
            j = 0;
            for (const auto& c : cams) { bad += f.camIds[j] != c.first || f.camT[3 * j] != c.second.t(0) || f.camQ[4 * j] != c.second.q(0); ++j; }
            std::printf("kept %zu of %zu mismatches %zu flatten_ms %.2f again_ms %.2f map_restatement_ms %.2f\n", n, obs.size(), bad, ms, ms2, ms_ref);
            return bad ? 1 : 0;
        }
        if (mode == "json") {   // vmm_host_test json <in.json> <tree_out.json> <stream_out.json>: the same document through both paths (no GPU)
            if (argc < 5) return 2;
            vmm_json::write_json(argv[3], vmm_json::read_json(dir));
            vmm_json::Writer w(argv[4]);
            // sections arrive in file order; a section without children is only known from the returned list
            struct Pending { std::string section, key; vmm_json::Node node; };
            std::vector<Pending> all;
            const std::vector<std::string> seen = vmm_json::read_json_sections(dir, [&](const std::string& section, const std::string& key, vmm_json::Node& n) {
                all.push_back(Pending{ section, key, std::move(n) });
            });
            size_t k = 0;
            for (const std::string& section : seen) {
                // the elements of this occurrence of the section: the run of `all` that starts at k with this name
                size_t e = k;
                while (e < all.size() && all[e].section == section) ++e;
                bool is_array = true;
                for (size_t i = k; i < e; ++i) is_array = is_array && all[i].key.empty();
                if (is_array) w.begin_array(section); else w.begin_object(section);
                for (size_t i = k; i < e; ++i) w.node(all[i].key, all[i].node);
                w.end();
                k = e;
            }
            w.close();
            return 0;
        }
        if (mode == "bigio") {   // vmm_host_test bigio <out_dir> <n tag observations>: streamed write + read of a large marker_detections.json (no GPU)
            if (argc < 4) return 2;
            const size_t n = (size_t)std::atoll(argv[3]), nImg = n / 40 + 1, nTag = n / 200 + 1;
            const std::string file = dir + "/marker_detections.json";
            struct rusage ru0;
            getrusage(RUSAGE_SELF, &ru0);   // the CUDA runtime the library links is resident before any I/O
            const auto tw0 = std::chrono::steady_clock::now();
            const auto coord = [](size_t i, int k) { return (double)(float)(3000.0 + 2500.0 * std::sin(0.001 * (double)i + k)); };
            {
                DetectionResult det;
                for (size_t i = 0; i < nImg; ++i) det.images.push_back(TagImg{ (int)i, dir + "/images/img_" + std::to_string(i) + ".jpg" });
                for (size_t i = 0; i < nTag; ++i) det.tags.push_back(Tag{ (int)i, "apriltag_36h11", 0.1285, 0.1295 });
                det.tagObservations.resize(n);
                for (size_t i = 0; i < n; ++i) {
                    TagObservation& o = det.tagObservations[i];
                    o.imageId = (int)(i / 40);
                    o.tagId = (int)(i % nTag);
                    for (int c = 0; c < 4; ++c) o.corners.push_back(Vector2d(coord(i, 2 * c), coord(i, 2 * c + 1)));
                }
                writeDetectionResult(det, file);
            }
            const auto tr0 = std::chrono::steady_clock::now();
            const DetectionResult back = readDetectionResult(file);
            const auto tr1 = std::chrono::steady_clock::now();
            size_t bad = (back.images.size() != nImg) + (back.tags.size() != nTag) + (back.tagObservations.size() != n);
            for (size_t i = 0; i < back.tagObservations.size() && !bad; ++i) {
                const TagObservation& o = back.tagObservations[i];
                bad += o.imageId != (int)(i / 40) || o.tagId != (int)(i % nTag) || o.corners.size() != 4;
                for (int c = 0; c < 4 && !bad; ++c) bad += o.corners[c].x() != coord(i, 2 * c) || o.corners[c].y() != coord(i, 2 * c + 1);
            }
            if (back.images.size() == nImg && nImg) bad += back.images[nImg - 1].filename != "images/img_" + std::to_string(nImg - 1) + ".jpg";
            struct rusage ru;
            getrusage(RUSAGE_SELF, &ru);
            std::ifstream in(file, std::ios::binary | std::ios::ate);
            std::printf("observations %zu mismatches %zu file_bytes %lld base_rss_kb %ld maxrss_kb %ld build_and_write_ms %lld read_ms %lld\n", n, bad, (long long)in.tellg(), ru0.ru_maxrss,
                ru.ru_maxrss, (long long)std::chrono::duration_cast<std::chrono::milliseconds>(tr0 - tw0).count(),
                (long long)std::chrono::duration_cast<std::chrono::milliseconds>(tr1 - tr0).count());
            return bad ? 1 : 0;
        }
        if (mode == "pnp") {
            std::ifstream in(dir);
            std::string word;
            while (in >> word) {
                if (word != "case") return 2;
                int n = 0, robust = 0;
                CameraModel m;
                in >> n >> robust >> m.fx >> m.fy >> m.cx >> m.cy;
                for (double& d : m.distortionCoefficients) in >> d;
                std::vector<Vector3d> X(n);
                std::vector<Vector2d> uv(n);
                for (int i = 0; i < n; ++i) in >> X[i](0) >> X[i](1) >> X[i](2) >> uv[i](0) >> uv[i](1);
                std::vector<unsigned char> inl(n, 1);
                const PnPResult r = robust ? solvePnPRobust(X, uv, m, 8.0, &inl) : solvePnP(X, uv, m);
                int cnt = 0;
                for (unsigned char c : inl) cnt += c;
                for (int k = 0; k < 9; ++k) std::printf("%.17g ", r.R.m[k]);
                std::printf("%.17g %.17g %.17g %.17g %d %d\n", r.t(0), r.t(1), r.t(2), r.rmsError, cnt, r.ok ? 1 : 0);
            }
            return 0;
        }
        if (mode == "moveorigin") {   // vmm_host_test moveorigin <project_dir> <tag id> <out.json>: no GPU
            if (argc < 5) return 2;
            DetectionResult det = readDetectionResult(dir + "/marker_detections.json");
            std::map<int, ReconstructedTag> tags;
            std::map<int, Camera> cams;
            CameraModel m = readCameraModel(dir + "/camera_intrinsics.json");
            parseReconstructions(dir + "/initial_reconstruction.json", tags, cams, m);
            TagReconstructor rec(std::move(det));
            rec.setCameraModel(m);
            rec.setQuiet(true);
            rec.setInitialReconstruction(tags, cams);
            rec.moveTagIntoOrigin(std::atoi(argv[3]));
            exportReconstructions(argv[4], rec.getReconstructedTags(), rec.getReconstructedCameras(), rec.getCameraModel());
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
