// TagReconstructor.cpp — the reference's TagReconstructor (src/TagReconstructor.cpp) as a thin host adapter:
// std::map state -> dense SoA arrays -> C ABI (include/vmm_ba.h) -> poses written back into the map nodes.
// What the reference delegates to ceres::Problem / ceres::Solve (:657-738) runs in libvmm_ba.so on the GPU.
#include "visual_marker_mapping/TagReconstructor.h"
#include "visual_marker_mapping/PnP.h"

#include "../../../../include/vmm_ba.h"

#include <limits>
#include <algorithm>
#include <cmath>
#include <functional>
#include <iostream>
#include <set>
#include <stdexcept>
#include <thread>

namespace visual_marker_mapping
{
std::vector<Vector3d> ReconstructedTag::computeLocalMarkerCorners3D() const
{
    std::vector<Vector3d> c;
    c.push_back(Vector3d(-tagWidth / 2.0, -tagHeight / 2.0, 0));
    c.push_back(Vector3d(tagWidth / 2.0, -tagHeight / 2.0, 0));
    c.push_back(Vector3d(tagWidth / 2.0, tagHeight / 2.0, 0));
    c.push_back(Vector3d(-tagWidth / 2.0, tagHeight / 2.0, 0));
    return c;
}

std::vector<Vector3d> ReconstructedTag::computeMarkerCorners3D() const
{
    std::vector<Vector3d> c = computeLocalMarkerCorners3D();
    const Matrix3d R = quaternionToRotationMatrix(q);
    for (size_t i = 0; i < 4; ++i) c[i] = R * c[i] + t;
    return c;
}

// dense view of the current maps (the problem assembly of src/TagReconstructor.cpp:663-724)
TagReconstructor::TagReconstructor(DetectionResult detection_result)
    : originTagId(-1), detectionResults_(std::move(detection_result))
{
}

TagReconstructor::~TagReconstructor()
{
    if (engine_) vmm_ba_destroy(engine_);
}

void TagReconstructor::setDevice(int cudaDevice)
{
    if (engine_) { vmm_ba_destroy(engine_); engine_ = nullptr; }
    device_ = cudaDevice;
}
void TagReconstructor::setParityMode(bool on) { parity_ = on; }
void TagReconstructor::setQuiet(bool on) { quiet_ = on; }

void TagReconstructor::moveTagIntoOrigin(int tagId)
{
    if (!reconstructedTags.count(tagId)) throw std::runtime_error("Tag with id " + std::to_string(tagId) + " is not reconstructed.");
    if (!quiet_) std::cout << "Transforming tag with id " << tagId << " into origin." << std::endl;
    // q = quat(tag)^-1 (Eigen's inverse: conjugate / squared norm), t = tag.t, :321-322
    const Vector4d qt = reconstructedTags[tagId].q;
    const double n2 = qt(0) * qt(0) + qt(1) * qt(1) + qt(2) * qt(2) + qt(3) * qt(3);
    const Vector4d q(qt(0) / n2, -qt(1) / n2, -qt(2) / n2, -qt(3) / n2);
    const Vector3d t = reconstructedTags[tagId].t;
    const Matrix3d R = quaternionToRotationMatrix(q);
    auto mul = [](const Vector4d& a, const Vector4d& b) {   // Hamilton product, (w,x,y,z)
        return Vector4d(a(0) * b(0) - a(1) * b(1) - a(2) * b(2) - a(3) * b(3), a(0) * b(1) + a(1) * b(0) + a(2) * b(3) - a(3) * b(2),
                        a(0) * b(2) - a(1) * b(3) + a(2) * b(0) + a(3) * b(1), a(0) * b(3) + a(1) * b(2) - a(2) * b(1) + a(3) * b(0));
    };
    for (auto& tag : reconstructedTags) {   // :324-328
        tag.second.q = mul(q, tag.second.q);
        tag.second.t = R * (tag.second.t - t);
    }
    for (auto& cam : reconstructedCameras) {   // :330-335 (the same two lines, although a camera pose maps world -> camera)
        cam.second.q = mul(q, cam.second.q);
        cam.second.t = R * (cam.second.t - t);
    }
    if (!quiet_) std::cout << "Finished transforming Tags" << std::endl;
}

int TagReconstructor::getLowestTag() const
{
    if (detectionResults_.tags.empty()) throw std::runtime_error("No tags in the detection result");   // reference: UB
    int lowest = detectionResults_.tags[0].tagId;
    for (const auto& tag : detectionResults_.tags) lowest = std::min(lowest, tag.tagId);
    return lowest;
}

const std::map<int, ReconstructedTag> TagReconstructor::getReconstructedTags() const { return reconstructedTags; }
const std::map<int, Camera> TagReconstructor::getReconstructedCameras() const { return reconstructedCameras; }
CameraModel TagReconstructor::getCameraModel() const { return camModel; }
void TagReconstructor::setCameraModel(const CameraModel& cameraModel) { camModel = cameraModel; }
void TagReconstructor::setOriginTagId(int id) { originTagId = id; }

void TagReconstructor::setInitialReconstruction(const std::map<int, ReconstructedTag>& tags, const std::map<int, Camera>& cameras)
{
    reconstructedTags = tags;
    reconstructedCameras = cameras;
    haveInitial_ = true;
}

void TagReconstructor::ensureEngine() const
{
    if (engine_) return;
    vmm_ba_handle h = nullptr;
    if (vmm_ba_create(device_, &h) != VMM_BA_OK) throw std::runtime_error(std::string("vmm_ba_create: ") + vmm_ba_last_error(nullptr));
    engine_ = h;
}

namespace
{
// id -> dense position for the ascending ids of a std::map: a direct table when the id range is compact (the usual
// 0..N-1 numbering of images and tags), else a binary search.  A site-sized problem looks up 8 M ids per flatten;
// std::map::find made that 2-6 s, longer than the bundle adjustment itself.
class IdIndex
{
public:
    explicit IdIndex(const std::vector<int>& ascendingIds) : ids_(ascendingIds)
    {
        if (ids_.empty()) return;
        lo_ = ids_.front();
        const long long range = (long long)ids_.back() - lo_ + 1;
        if (range <= 8 * (long long)ids_.size() + 1024) {
            table_.assign((size_t)range, -1);
            for (size_t i = 0; i < ids_.size(); ++i) table_[(size_t)(ids_[i] - lo_)] = (int)i;
        }
    }
    int find(int id) const   // -1: not reconstructed
    {
        if (!table_.empty()) {
            const long long o = (long long)id - lo_;
            return (o < 0 || o >= (long long)table_.size()) ? -1 : table_[(size_t)o];
        }
        const auto it = std::lower_bound(ids_.begin(), ids_.end(), id);
        return (it == ids_.end() || *it != id) ? -1 : (int)(it - ids_.begin());
    }

private:
    const std::vector<int>& ids_;
    std::vector<int> table_;
    int lo_ = 0;
};
}

void TagReconstructor::flatten(Flat& f) const
{
    // f may be the buffer of an earlier call (the class keeps one): the pose arrays are rebuilt, the per-observation
    // arrays only resized — no zero fill and no page faults for the 250 MB of a site-sized problem
    f.camIds.clear(); f.camQ.clear(); f.camT.clear();
    f.tagIds.clear(); f.tagQ.clear(); f.tagT.clear(); f.tagWH.clear();
    f.fixedTag = -1;
    f.camIds.reserve(reconstructedCameras.size());
    f.camQ.reserve(4 * reconstructedCameras.size());
    f.camT.reserve(3 * reconstructedCameras.size());
    for (const auto& it : reconstructedCameras) {
        f.camIds.push_back(it.first);
        for (int k = 0; k < 4; ++k) f.camQ.push_back(it.second.q(k));
        for (int k = 0; k < 3; ++k) f.camT.push_back(it.second.t(k));
    }
    f.tagIds.reserve(reconstructedTags.size());
    for (const auto& it : reconstructedTags) {
        if (it.first == originTagId) f.fixedTag = (long long)f.tagIds.size();   // SetParameterBlockConstant, :669-673
        f.tagIds.push_back(it.first);
        for (int k = 0; k < 4; ++k) f.tagQ.push_back(it.second.q(k));
        for (int k = 0; k < 3; ++k) f.tagT.push_back(it.second.t(k));
        f.tagWH.push_back(it.second.tagWidth);    // corners come from the RECONSTRUCTED tag's size (:713)
        f.tagWH.push_back(it.second.tagHeight);
    }
    const IdIndex camIndex(f.camIds), tagIndex(f.tagIds);
    // the tag observations whose image and tag are both reconstructed (:699-724), in their order: chunks of the list
    // are classified and counted by one thread each, then written at their prefix offsets
    const auto& obs = detectionResults_.tagObservations;
    const size_t n = obs.size();
    const unsigned hw = std::thread::hardware_concurrency();
    const size_t T = n >= ((size_t)1 << 16) ? std::max<size_t>(1, std::min<size_t>(hw ? hw : 1, 16)) : 1;
    std::vector<size_t> kept(T + 1, 0), badAt(T, n);
    std::vector<int32_t>& cam = flatScratchCam_;
    std::vector<int32_t>& tag = flatScratchTag_;
    if (cam.size() < n) { cam.resize(n); tag.resize(n); }
    const auto chunk = [&](size_t c) { return std::make_pair(n * c / T, n * (c + 1) / T); };
    const auto run = [&](const std::function<void(size_t)>& body) {
        if (T == 1) { body(0); return; }
        std::vector<std::thread> th;
        for (size_t c = 0; c < T; ++c) th.emplace_back(body, c);
        for (auto& t : th) t.join();
    };
    run([&](size_t c) {
        size_t cnt = 0;
        for (size_t k = chunk(c).first; k < chunk(c).second; ++k) {
            const int ci = camIndex.find(obs[k].imageId);
            const int ti = ci < 0 ? -1 : tagIndex.find(obs[k].tagId);
            cam[k] = ti < 0 ? -1 : ci;
            tag[k] = ti;
            if (ti < 0) continue;
            if (obs[k].corners.size() != 4 && badAt[c] == n) badAt[c] = k;
            ++cnt;
        }
        kept[c + 1] = cnt;
    });
    for (size_t c = 0; c < T; ++c)
        if (badAt[c] != n) throw std::runtime_error("Unexpected number of values");
    for (size_t c = 0; c < T; ++c) kept[c + 1] += kept[c];
    const size_t m = kept[T];
    f.toCam.resize(m);
    f.toTag.resize(m);
    f.toUV.resize(8 * m);
    f.toSource.resize(m);
    run([&](size_t c) {
        size_t o = kept[c];
        for (size_t k = chunk(c).first; k < chunk(c).second; ++k) {
            if (cam[k] < 0) continue;
            f.toCam[o] = cam[k];
            f.toTag[o] = tag[k];
            for (int i = 0; i < 4; ++i) { f.toUV[8 * o + 2 * i] = obs[k].corners[i].x(); f.toUV[8 * o + 2 * i + 1] = obs[k].corners[i].y(); }
            f.toSource[o] = k;
            ++o;
        }
    });
}

static void check(vmm_ba_engine* e, int rc, const char* what)
{
    if (rc != VMM_BA_OK) throw std::runtime_error(std::string(what) + ": " + vmm_ba_last_error(e));
}

static void upload(vmm_ba_engine* e, const CameraModel& m, const std::vector<double>& camQ, const std::vector<double>& camT,
    const std::vector<double>& tagQ, const std::vector<double>& tagT, const std::vector<double>& tagWH, long long fixedTag,
    const std::vector<int32_t>& toCam, const std::vector<int32_t>& toTag, const std::vector<double>& toUV)
{
    check(e, vmm_ba_set_intrinsics(e, m.fx, m.fy, m.cx, m.cy, m.distortionCoefficients), "vmm_ba_set_intrinsics");
    check(e, vmm_ba_set_problem(e, (int64_t)(camQ.size() / 4), camQ.data(), camT.data(), (int64_t)(tagQ.size() / 4), tagQ.data(),
                 tagT.data(), tagWH.data(), fixedTag, (int64_t)toCam.size(), toCam.data(), toTag.data(), toUV.data()),
        "vmm_ba_set_problem");
}

void TagReconstructor::doBundleAdjustment(int maxNumIterations, size_t /*ceresThreads*/, bool robustify, bool printSummary)
{
    if (reconstructedTags.empty()) throw std::runtime_error("doBundleAdjustment: no reconstructed tags");
    Flat& f = flat_;
    flatten(f);
    ensureEngine();
    upload(engine_, camModel, f.camQ, f.camT, f.tagQ, f.tagT, f.tagWH, f.fixedTag, f.toCam, f.toTag, f.toUV);

    vmm_ba_options opt;
    vmm_ba_options_init(&opt);
    opt.max_num_iterations = maxNumIterations;                                 // :731
    opt.loss_kind = robustify ? VMM_BA_LOSS_HUBER : VMM_BA_LOSS_NONE;          // new ceres::HuberLoss(1.0), :721
    opt.loss_scale = 1.0;
    opt.parity_mode = parity_ ? 1 : 0;
    vmm_ba_summary s;
    check(engine_, vmm_ba_solve(engine_, &opt, &s), "vmm_ba_solve");
    check(engine_, vmm_ba_get_poses(engine_, f.camQ.data(), f.camT.data(), f.tagQ.data(), f.tagT.data()), "vmm_ba_get_poses");

    // in-place result, as Ceres writes through the pointers into the map nodes (:665-666, :692-693)
    for (size_t i = 0; i < f.camIds.size(); ++i) {
        Camera& c = reconstructedCameras[f.camIds[i]];
        for (int k = 0; k < 4; ++k) c.q(k) = f.camQ[4 * i + k];
        for (int k = 0; k < 3; ++k) c.t(k) = f.camT[3 * i + k];
    }
    for (size_t j = 0; j < f.tagIds.size(); ++j) {
        ReconstructedTag& t = reconstructedTags[f.tagIds[j]];
        for (int k = 0; k < 4; ++k) t.q(k) = f.tagQ[4 * j + k];
        for (int k = 0; k < 3; ++k) t.t(k) = f.tagT[3 * j + k];
    }
    report_.terminationType = s.termination_type;
    report_.initialCost = s.initial_cost;
    report_.finalCost = s.final_cost;
    report_.numSuccessfulSteps = s.num_successful_steps;
    report_.numUnsuccessfulSteps = s.num_unsuccessful_steps;
    report_.numIterations = s.num_iterations;
    report_.pcgIterations = s.pcg_iterations;
    report_.kernelLaunches = s.kernel_launches;
    report_.timeTotal = s.time_total_s;
    report_.timeLinearize = s.time_linearize_s;
    report_.timeLinearSolver = s.time_linear_solver_s;
    report_.timeCostEval = s.time_cost_eval_s;
    report_.message = s.message;
    if (!quiet_) std::cout << "Solution " << s.termination_type << std::endl;   // :740 (never acted on)
    if (printSummary && !quiet_) {
        std::cout << "vmm_ba report: " << f.camIds.size() << " cameras, " << f.tagIds.size() << " tags, " << f.toCam.size() * 4
                  << " residual blocks\n  cost " << s.initial_cost << " -> " << s.final_cost << "\n  iterations " << s.num_iterations
                  << " (successful " << s.num_successful_steps << ", unsuccessful " << s.num_unsuccessful_steps << "), PCG iterations "
                  << s.pcg_iterations << "\n  time total " << s.time_total_s << " s (linearize " << s.time_linearize_s << ", linear solver "
                  << s.time_linear_solver_s << ", cost evaluation " << s.time_cost_eval_s << ")\n  termination: " << s.message << std::endl;
    }
    // :744-783 — standard deviations of the tag translations from the covariance blocks, same lines as the reference
    // ("StdDevNorm" is the square root of the norm of the standard deviations, :772-773, reproduced as it is)
    if (printSummary) {
        // nothing of an earlier call may be read as this one's result
        tagCovariances_.clear();
        markerPositionRms_ = std::numeric_limits<double>::quiet_NaN();
        std::vector<double> cov(9 * f.tagIds.size(), 0.0);
        const int rc = vmm_ba_marker_translation_covariance(engine_, opt.loss_kind, opt.loss_scale, cov.data());
        if (rc != VMM_BA_OK) {
            if (!quiet_) std::cout << "Covariance not available: " << vmm_ba_last_error(engine_) << std::endl;
            return;
        }
        double avg[3] = { 0.0, 0.0, 0.0 };
        for (size_t j = 0; j < f.tagIds.size(); ++j) {
            const double* c = cov.data() + 9 * j;
            tagCovariances_[f.tagIds[j]] = { c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7], c[8] };
            const double sx = std::sqrt(c[0]), sy = std::sqrt(c[4]), sz = std::sqrt(c[8]);
            if (!quiet_)
                std::cout << "StdDev of tag " << f.tagIds[j] << ": " << sx << " " << sy << " " << sz
                          << " | StdDevNorm: " << std::sqrt(std::sqrt(sx * sx + sy * sy + sz * sz)) << std::endl;
            avg[0] += c[0]; avg[1] += c[4]; avg[2] += c[8];
        }
        for (double& a : avg) a /= (double)f.tagIds.size();
        markerPositionRms_ = std::sqrt(avg[0] + avg[1] + avg[2]);   // norm of the per-axis RMS standard deviations, :777-781
        if (!quiet_) std::cout << "Marker Position RMS = " << markerPositionRms_ << std::endl;
    }
}

std::vector<double> TagReconstructor::reprojectionErrorSums(const Flat& f) const
{
    std::vector<double> sums(f.toCam.size(), 0.0);
    if (f.toCam.empty() || f.tagIds.empty()) return sums;
    ensureEngine();
    upload(engine_, camModel, f.camQ, f.camT, f.tagQ, f.tagT, f.tagWH, f.fixedTag, f.toCam, f.toTag, f.toUV);
    check(engine_, vmm_ba_reprojection_error_sums(engine_, sums.data()), "vmm_ba_reprojection_error_sums");
    return sums;
}

namespace
{
// Per-block sums of the per-observation error sums, accumulated in observation order (the order the reference's
// std::map accumulators see, so the doubles are the same) in dense arrays instead of 4 M map look-ups.
void accumulateByBlock(const std::vector<int32_t>& block, const std::vector<double>& sums, size_t numBlocks, std::vector<double>& err,
    std::vector<long long>& num)
{
    err.assign(numBlocks, 0.0);
    num.assign(numBlocks, 0);
    for (size_t k = 0; k < sums.size(); ++k) {
        num[(size_t)block[k]] += 4;
        err[(size_t)block[k]] += sums[k];
    }
}
}

const std::map<int, double> TagReconstructor::computeReprojectionErrorPerImg() const
{
    Flat& f = flat_;
    flatten(f);
    const std::vector<double> sums = reprojectionErrorSums(f);
    std::vector<double> e;
    std::vector<long long> n;
    accumulateByBlock(f.toCam, sums, f.camIds.size(), e, n);
    std::map<int, double> err;
    for (size_t i = 0; i < f.camIds.size(); ++i)     // cameras without observations: -1 (:379-383)
        err.emplace_hint(err.end(), f.camIds[i], n[i] ? e[i] / (double)n[i] : -1.0);
    return err;
}

const std::map<int, double> TagReconstructor::computeReprojectionErrorPerTag(double& avg) const
{
    Flat& f = flat_;
    flatten(f);
    const std::vector<double> sums = reprojectionErrorSums(f);
    std::vector<double> e;
    std::vector<long long> n;
    accumulateByBlock(f.toTag, sums, f.tagIds.size(), e, n);
    std::map<int, double> err;
    avg = 0;
    long long total = 0;
    for (size_t j = 0; j < f.tagIds.size(); ++j) {   // avg = sum of the per-tag sums / total corner count (:416-426)
        if (!n[j]) continue;                         // a tag without observations has no entry
        avg += e[j];
        total += n[j];
        err.emplace_hint(err.end(), f.tagIds[j], e[j] / (double)n[j]);
    }
    if (total) avg /= (double)total;
    return err;
}

const std::vector<Vector2d> TagReconstructor::computeReprojectionErrorPerCorner() const
{
    // host loop like the reference (:430-455); it is not called by either tool
    std::vector<Vector2d> out;
    for (const auto& obs : detectionResults_.tagObservations) {
        const auto c = reconstructedCameras.find(obs.imageId);
        const auto t = reconstructedTags.find(obs.tagId);
        if (c == reconstructedCameras.end() || t == reconstructedTags.end()) continue;
        const std::vector<Vector3d> pts = t->second.computeMarkerCorners3D();
        const Matrix3d R = quaternionToRotationMatrix(c->second.q);
        for (size_t i = 0; i < pts.size(); ++i) out.push_back(camModel.projectPoint(R * pts[i] + c->second.t) - obs.corners[i]);
    }
    return out;
}

void TagReconstructor::removeBadMarkers(double threshold)
{
    double avg = 0;
    const auto errs = computeReprojectionErrorPerTag(avg);
    for (const auto& tr : errs) {
        if (tr.second > threshold && tr.first != originTagId) {
            if (!quiet_) std::cout << "Removing bad marker with id " << tr.first << " and reprojection error " << tr.second << std::endl;
            reconstructedTags.erase(tr.first);
        }
    }
}

void TagReconstructor::removeBadCameras(double threshold)
{
    const auto errs = computeReprojectionErrorPerImg();
    for (const auto& tr : errs) {
        if (tr.second > threshold || tr.second < 0) {
            if (!quiet_) std::cout << "Removing bad camera with id " << tr.first << " and reprojection error " << tr.second << std::endl;
            reconstructedCameras.erase(tr.first);
        }
    }
}

bool TagReconstructor::computeRelativeCameraPoseFromImg(int imageId, Vector4d& q, Vector3d& t)
{
    // :280-312 — all corners of already reconstructed tags seen in this image against their current world positions
    std::vector<Vector3d> markerCorners3D;
    std::vector<Vector2d> observations2D;
    for (const auto& tagObs : detectionResults_.tagObservations) {
        if (tagObs.imageId != imageId) continue;
        const auto tagIt = reconstructedTags.find(tagObs.tagId);
        if (tagIt == reconstructedTags.end()) continue;
        const std::vector<Vector3d> tagCorners = tagIt->second.computeMarkerCorners3D();
        markerCorners3D.insert(markerCorners3D.begin(), tagCorners.begin(), tagCorners.end());
        observations2D.insert(observations2D.begin(), tagObs.corners.begin(), tagObs.corners.end());
    }
    if (!quiet_) std::cout << "   Reconstructing camera pose from " << observations2D.size() << " 2d/3d correspondences" << std::endl;
    if (observations2D.empty()) return false;
    const PnPResult r = solvePnPRobust(markerCorners3D, observations2D, camModel);   // replaces solvePnPRansacEigen, :306
    if (!r.ok) throw std::runtime_error("Camera pose of image " + std::to_string(imageId) + " could not be initialised (degenerate correspondences)");
    q = rotationMatrixToQuaternion(r.R);
    t = r.t;
    return true;
}

void TagReconstructor::initialiseTagsSeenFrom(int imageId, const std::map<int, std::set<int> >& whichImagesObserveTag,
                                              const std::map<int, const Tag*>& tagById)
{
    // :180-240 — tags seen in this image that are not reconstructed yet: T_world<-tag = T_cam<-world^-1 * T_cam<-tag
    const Camera& cam = reconstructedCameras.at(imageId);
    const Matrix3d Rc = quaternionToRotationMatrix(cam.q);
    for (const auto& tagObs : detectionResults_.tagObservations) {
        if (tagObs.imageId != imageId) continue;
        if (reconstructedTags.count(tagObs.tagId)) continue;
        const auto seen = whichImagesObserveTag.find(tagObs.tagId);
        if (seen == whichImagesObserveTag.end() || seen->second.size() < 2) {
            if (!quiet_) std::cout << "   Skipping reconstruction of tag " << tagObs.tagId << ": Only observed once!" << std::endl;
            continue;
        }
        const auto tg = tagById.find(tagObs.tagId);
        if (tg == tagById.end()) continue;   // the reference dereferences a null pointer here (tagById[...]->tagType, :211)
        ReconstructedTag recTag;
        recTag.id = tagObs.tagId;
        recTag.tagType = tg->second->tagType;
        recTag.tagWidth = tg->second->width;
        recTag.tagHeight = tg->second->height;
        const PnPResult r = solvePnP(recTag.computeLocalMarkerCorners3D(), tagObs.corners, camModel);   // replaces solvePnPEigen, :221
        if (!r.ok) {
            if (!quiet_) std::cout << "   Skipping reconstruction of tag " << tagObs.tagId << ": degenerate corner configuration" << std::endl;
            continue;
        }
        // extrinsic^-1 * T: R_w = Rc^T R, t_w = Rc^T (t - tc)
        Matrix3d Rw;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Rw.m[i * 3 + j] = Rc.m[i] * r.R.m[j] + Rc.m[3 + i] * r.R.m[3 + j] + Rc.m[6 + i] * r.R.m[6 + j];
        const Vector3d d = r.t - cam.t;
        recTag.q = rotationMatrixToQuaternion(Rw);
        recTag.t = Vector3d(Rc.m[0] * d(0) + Rc.m[3] * d(1) + Rc.m[6] * d(2), Rc.m[1] * d(0) + Rc.m[4] * d(1) + Rc.m[7] * d(2),
                            Rc.m[2] * d(0) + Rc.m[5] * d(1) + Rc.m[8] * d(2));
        reconstructedTags.emplace(tagObs.tagId, recTag);
        if (!quiet_) std::cout << "   Initialized tag with id " << tagObs.tagId << std::endl;
    }
}

void TagReconstructor::startReconstruction(size_t numThreads)
{
    if (originTagId == -1) originTagId = getLowestTag();   // :89-92
    if (haveInitial_) {
        // initial poses were handed in (setInitialReconstruction): only the closing sequence of the reference, :271-277
        if (reconstructedTags.find(originTagId) == reconstructedTags.end())
            throw std::runtime_error("Origin tag " + std::to_string(originTagId) + " is not among the initial tags");
        if (!quiet_) std::cout << "Start bundle adjustment with " << reconstructedCameras.size() << " cameras and " << reconstructedTags.size() << " tags" << std::endl;
        doBundleAdjustment(1500, numThreads, true, false);
        removeBadMarkers(2.0);
        removeBadCameras(2.0);
        doBundleAdjustment(1500, numThreads, false, true);
        return;
    }
    // ---- the incremental initialisation of the reference, src/TagReconstructor.cpp:97-267 ----
    std::map<int, std::set<int> > whichImagesObserveTag, tagsInImage;
    for (const auto& tagObs : detectionResults_.tagObservations) {
        whichImagesObserveTag[tagObs.tagId].insert(tagObs.imageId);
        tagsInImage[tagObs.imageId].insert(tagObs.tagId);
    }
    std::map<int, std::string> imageFilenames;
    for (const auto& img : detectionResults_.images) imageFilenames[img.imageId] = img.filename;
    std::map<int, const Tag*> tagById;
    for (const auto& tag : detectionResults_.tags) tagById[tag.tagId] = &tag;

    // the image with the most tags among those that see the origin tag, :113-124
    size_t maxObservations = 0;
    int curImageId = -1;
    for (int imageId : whichImagesObserveTag[originTagId])
        if (tagsInImage[imageId].size() > maxObservations) { curImageId = imageId; maxObservations = tagsInImage[imageId].size(); }
    if (tagById.find(originTagId) == tagById.end())
        throw std::runtime_error("Could not use tag with id " + std::to_string(originTagId) + " as origin tag, because it was not detected.");
    if (curImageId == -1)
        throw std::runtime_error("Could not use tag with id " + std::to_string(originTagId) + " as origin tag, because no image observes it.");

    reconstructedTags.clear();
    reconstructedCameras.clear();
    ReconstructedTag origin;   // :134-141
    origin.id = originTagId;
    origin.tagType = tagById[originTagId]->tagType;
    origin.tagWidth = tagById[originTagId]->width;
    origin.tagHeight = tagById[originTagId]->height;
    reconstructedTags.emplace(originTagId, origin);

    while (true) {
        // one camera per round (:146-267); in wave mode every image that already sees a reconstructed tag joins the
        // round, so that the number of bundle adjustments grows with the depth of the map, not with its images
        std::vector<int> round{ curImageId };
        if (waveMode_) {
            round.clear();
            for (const auto& img : detectionResults_.images) {
                if (reconstructedCameras.count(img.imageId)) continue;
                bool sees = false;
                for (int tagId : tagsInImage[img.imageId]) sees = sees || reconstructedTags.count(tagId);
                if (sees) round.push_back(img.imageId);
            }
        }
        for (int imageId : round) {
            if (!quiet_)
                std::cout << "Reconstructing image " << reconstructedCameras.size() << "/" << detectionResults_.images.size() << " | "
                          << imageFilenames[imageId] << " with id: " << imageId << std::endl;
            Camera newCamera;
            newCamera.cameraId = imageId;
            if (!computeRelativeCameraPoseFromImg(imageId, newCamera.q, newCamera.t))
                throw std::runtime_error("No reconstructed tags in image found. To reconstruct the image pose "
                                         "already reconstructed markers are needed. This should NOT happen.");
            if (!quiet_) std::cout << "   Initialized camera with id " << imageId << std::endl;
            reconstructedCameras.emplace(imageId, newCamera);
        }
        for (int imageId : round) initialiseTagsSeenFrom(imageId, whichImagesObserveTag, tagById);

        doBundleAdjustment(400, numThreads, true);   // :243
        removeBadMarkers(2.0);                       // :246

        // next image: the unreconstructed one with the most reconstructed tags, :248-263
        int maxPairs = 0;
        for (const auto& img : detectionResults_.images) {
            if (reconstructedCameras.count(img.imageId)) continue;
            int numCorrespondences = 0;
            for (int tagId : tagsInImage[img.imageId])
                if (reconstructedTags.count(tagId)) numCorrespondences++;
            if (numCorrespondences > maxPairs) { maxPairs = numCorrespondences; curImageId = img.imageId; }
        }
        if (maxPairs == 0) break;
        if (!quiet_) std::cout << "----------------------------------------------------------" << std::endl;
    }
    if (!quiet_) std::cout << "Starting final bundle adjustment" << std::endl;
    doBundleAdjustment(1500, numThreads, true, false);   // :271-277
    removeBadMarkers(2.0);
    removeBadCameras(2.0);
    doBundleAdjustment(1500, numThreads, false, true);
}
}
