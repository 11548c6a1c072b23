// TagReconstructor.cpp — the reference's TagReconstructor (src/TagReconstructor.cpp) as a thin host adapter:
// std::map state -> dense SoA arrays -> C ABI (include/vmm_ba.h) -> poses written back into the map nodes.
// What the reference delegates to ceres::Problem / ceres::Solve (:657-738) runs in libvmm_ba.so on the GPU.
#include "visual_marker_mapping/TagReconstructor.h"

#include "../../../../include/vmm_ba.h"

#include <algorithm>
#include <cmath>
#include <iostream>
#include <stdexcept>

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
struct TagReconstructor::Flat
{
    std::vector<int> camIds, tagIds;                  // dense index -> id (ascending, std::map order)
    std::vector<double> camQ, camT, tagQ, tagT, tagWH;
    std::vector<int32_t> toCam, toTag;
    std::vector<double> toUV;
    std::vector<size_t> toSource;                     // index into detectionResults_.tagObservations
    long long fixedTag = -1;
};

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

void TagReconstructor::flatten(Flat& f) const
{
    std::map<int, int> camIndex, tagIndex;
    for (const auto& it : reconstructedCameras) {
        camIndex[it.first] = (int)f.camIds.size();
        f.camIds.push_back(it.first);
        for (int k = 0; k < 4; ++k) f.camQ.push_back(it.second.q(k));
        for (int k = 0; k < 3; ++k) f.camT.push_back(it.second.t(k));
    }
    for (const auto& it : reconstructedTags) {
        if (it.first == originTagId) f.fixedTag = (long long)f.tagIds.size();   // SetParameterBlockConstant, :669-673
        tagIndex[it.first] = (int)f.tagIds.size();
        f.tagIds.push_back(it.first);
        for (int k = 0; k < 4; ++k) f.tagQ.push_back(it.second.q(k));
        for (int k = 0; k < 3; ++k) f.tagT.push_back(it.second.t(k));
        f.tagWH.push_back(it.second.tagWidth);    // corners come from the RECONSTRUCTED tag's size (:713)
        f.tagWH.push_back(it.second.tagHeight);
    }
    const auto& obs = detectionResults_.tagObservations;
    for (size_t k = 0; k < obs.size(); ++k) {
        const auto c = camIndex.find(obs[k].imageId);
        if (c == camIndex.end()) continue;
        const auto t = tagIndex.find(obs[k].tagId);
        if (t == tagIndex.end()) continue;
        if (obs[k].corners.size() != 4) throw std::runtime_error("Unexpected number of values");
        f.toCam.push_back(c->second);
        f.toTag.push_back(t->second);
        for (int i = 0; i < 4; ++i) { f.toUV.push_back(obs[k].corners[i].x()); f.toUV.push_back(obs[k].corners[i].y()); }
        f.toSource.push_back(k);
    }
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
    Flat f;
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
        std::vector<double> cov(9 * f.tagIds.size(), 0.0);
        const int rc = vmm_ba_marker_translation_covariance(engine_, opt.loss_kind, opt.loss_scale, cov.data());
        if (rc != VMM_BA_OK) {
            if (!quiet_) std::cout << "Covariance not available: " << vmm_ba_last_error(engine_) << std::endl;
            return;
        }
        tagCovariances_.clear();
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

const std::map<int, double> TagReconstructor::computeReprojectionErrorPerImg() const
{
    Flat f;
    flatten(f);
    const std::vector<double> sums = reprojectionErrorSums(f);
    std::map<int, double> err;
    std::map<int, int> num;
    for (size_t k = 0; k < sums.size(); ++k) {
        const int id = f.camIds[f.toCam[k]];
        num[id] += 4;
        err[id] += sums[k];
    }
    for (auto& e : err) e.second /= num[e.first];
    for (const auto& cam : reconstructedCameras)     // cameras without observations: -1 (:379-383)
        if (err.count(cam.first) == 0) err[cam.first] = -1.0;
    return err;
}

const std::map<int, double> TagReconstructor::computeReprojectionErrorPerTag(double& avg) const
{
    Flat f;
    flatten(f);
    const std::vector<double> sums = reprojectionErrorSums(f);
    std::map<int, double> err;
    std::map<int, int> num;
    for (size_t k = 0; k < sums.size(); ++k) {
        const int id = f.tagIds[f.toTag[k]];
        num[id] += 4;
        err[id] += sums[k];
    }
    avg = 0;
    int total = 0;
    for (auto& e : err) {       // avg = sum of the per-tag sums / total corner count (:416-426)
        avg += e.second;
        e.second /= num[e.first];
        total += num[e.first];
    }
    if (total) avg /= total;
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

void TagReconstructor::startReconstruction(size_t numThreads)
{
    if (originTagId == -1) originTagId = getLowestTag();   // :89-92
    if (!haveInitial_)
        throw std::runtime_error("startReconstruction: the incremental PnP initialisation (src/TagReconstructor.cpp:146-267) is not "
                                 "part of this build; provide initial poses with setInitialReconstruction()");
    if (reconstructedTags.find(originTagId) == reconstructedTags.end())
        throw std::runtime_error("Origin tag " + std::to_string(originTagId) + " is not among the initial tags");
    if (!quiet_) std::cout << "Start bundle adjustment with " << reconstructedCameras.size() << " cameras and " << reconstructedTags.size() << " tags" << std::endl;
    // the closing sequence of the reference, :271-277
    doBundleAdjustment(1500, numThreads, true, false);
    removeBadMarkers(2.0);
    removeBadCameras(2.0);
    doBundleAdjustment(1500, numThreads, false, true);
}
}
