// TagReconstructor.h — the reference's public class (include/visual_marker_mapping/TagReconstructor.h:58-146)
// over the B200 engine.  Same members, same argument meaning, same error behaviour; the Ceres problem
// behind doBundleAdjustment (src/TagReconstructor.cpp:646-784) is replaced by libvmm_ba.so through the C ABI
// in include/vmm_ba.h.  There is no CPU fallback: without a CUDA device every solving member throws.
#ifndef VISUAL_MARKER_MAPPING_TAGRECONSTRUCTOR_H_
#define VISUAL_MARKER_MAPPING_TAGRECONSTRUCTOR_H_

#include "visual_marker_mapping/Camera.h"
#include "visual_marker_mapping/CameraModel.h"
#include "visual_marker_mapping/DetectionResults.h"
#include <array>
#include <cstddef>
#include <cstdint>
#include <map>
#include <set>
#include <string>
#include <vector>

struct vmm_ba_engine;

namespace visual_marker_mapping
{
struct ReconstructedTag
{
    int id = -1;
    std::string tagType = "";
    Vector4d q = Vector4d::UnitX();   // marker -> world, (w,x,y,z)
    Vector3d t = Vector3d::Zero();
    double tagWidth = 0.0;
    double tagHeight = 0.0;

    std::vector<Vector3d> computeLocalMarkerCorners3D() const;   // TagReconstructor.h:44-52
    std::vector<Vector3d> computeMarkerCorners3D() const;        // TagReconstructor.h:33-43
};

// what the engine reported for the last doBundleAdjustment (the reference only prints it)
struct BundleAdjustmentReport
{
    int terminationType = -1;
    double initialCost = 0, finalCost = 0;
    int numSuccessfulSteps = 0, numUnsuccessfulSteps = 0, numIterations = 0;
    long long pcgIterations = 0, kernelLaunches = 0;
    double timeTotal = 0, timeLinearize = 0, timeLinearSolver = 0, timeCostEval = 0;
    std::string message;
};

class TagReconstructor
{
public:
    explicit TagReconstructor(DetectionResult detection_result);
    ~TagReconstructor();
    TagReconstructor(const TagReconstructor&) = delete;
    TagReconstructor& operator=(const TagReconstructor&) = delete;

    // Reference: incremental PnP initialisation + BA per camera, then BA / cull / BA
    // (src/TagReconstructor.cpp:86-278), with the in-repo PnP of PnP.h in place of the OpenCV calls.
    // When initial poses were given with setInitialReconstruction() only the closing sequence (:271-277) runs.
    void startReconstruction(size_t numThreads = 1);

    // initial camera / tag poses (e.g. parseReconstructions of an earlier run)
    void setInitialReconstruction(const std::map<int, ReconstructedTag>& tags, const std::map<int, Camera>& cameras);

    const std::map<int, double> computeReprojectionErrorPerImg() const;
    const std::map<int, double> computeReprojectionErrorPerTag(double& avg) const;
    const std::vector<Vector2d> computeReprojectionErrorPerCorner() const;

    // TagReconstructor.h:110 / src/TagReconstructor.cpp:314-338: re-express every pose relative to tag `tagId`
    // (the reference's formulas, applied to tags and cameras alike, reproduced as they are)
    void moveTagIntoOrigin(int tagId);

    int getLowestTag() const;
    const std::map<int, ReconstructedTag> getReconstructedTags() const;
    const std::map<int, Camera> getReconstructedCameras() const;

    void removeBadMarkers(double threshold);
    void removeBadCameras(double threshold);

    CameraModel getCameraModel() const;
    void setCameraModel(const CameraModel& cameraModel);
    void setOriginTagId(int originTagId);

    // the protected seam of the reference (TagReconstructor.h:131-132), public here so that callers
    // that only want the refinement can drive it; ceresThreads is accepted and ignored.
    void doBundleAdjustment(int maxNumIterations, size_t ceresThreads, bool robustify = true, bool printSummary = false);

    // extensions
    void setDevice(int cudaDevice);
    void setParityMode(bool on);     // true: PCG to 1e-13 so that iterates track an exact-solve (Ceres) trajectory
    void setQuiet(bool on);
    // true: every image that already sees a reconstructed tag is initialised in the same round and the bundle
    // adjustment runs once per round instead of once per image; false = the reference's schedule
    void setWaveMode(bool on) { waveMode_ = on; }
    const BundleAdjustmentReport& lastReport() const { return report_; }
    // what the reference only prints (src/TagReconstructor.cpp:744-783): 3x3 translation covariance per tag id
    // (row-major) and the "Marker Position RMS" of the last doBundleAdjustment(..., printSummary = true)
    const std::map<int, std::array<double, 9>>& tagTranslationCovariances() const { return tagCovariances_; }
    double markerPositionRms() const { return markerPositionRms_; }

    // the problem as doBundleAdjustment hands it to vmm_ba_set_problem (src/TagReconstructor.cpp:663-724 flattened):
    // dense 0-based positions in std::map order, ids stay on the host side.  No GPU needed.
    struct FlatProblem
    {
        std::vector<int> camIds, tagIds;                  // dense index -> id (ascending, std::map order)
        std::vector<double> camQ, camT, tagQ, tagT, tagWH;
        std::vector<int32_t> toCam, toTag;
        std::vector<double> toUV;
        std::vector<size_t> toSource;                     // index into the detection result's tagObservations
        long long fixedTag = -1;
    };
    const FlatProblem& flatProblem() const   // valid until the next call that flattens (this one, doBundleAdjustment, the statistics)
    {
        flatten(flat_);
        return flat_;
    }

protected:
    typedef FlatProblem Flat;
    void flatten(Flat& f) const;
    void ensureEngine() const;
    std::vector<double> reprojectionErrorSums(const Flat& f) const;
    bool computeRelativeCameraPoseFromImg(int imageId, Vector4d& q, Vector3d& t);   // TagReconstructor.h:134-136
    void initialiseTagsSeenFrom(int imageId, const std::map<int, std::set<int> >& whichImagesObserveTag, const std::map<int, const Tag*>& tagById);

    int originTagId;
    DetectionResult detectionResults_;
    std::map<int, ReconstructedTag> reconstructedTags;
    std::map<int, Camera> reconstructedCameras;
    CameraModel camModel;

    mutable vmm_ba_engine* engine_ = nullptr;
    mutable Flat flat_;                                           // reused by every doBundleAdjustment / statistics call
    mutable std::vector<int32_t> flatScratchCam_, flatScratchTag_;
    int device_ = 0;
    bool parity_ = false, quiet_ = false, haveInitial_ = false, waveMode_ = false;
    BundleAdjustmentReport report_;
    std::map<int, std::array<double, 9>> tagCovariances_;
    double markerPositionRms_ = 0.0;
};
}
#endif
