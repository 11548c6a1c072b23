// PnP.h — pose from 2D/3D correspondences for the incremental initialisation of the map.
//
// Replaces the two OpenCV wrappers of the reference (src/EigenCVConversions.cpp:38-106):
//   solvePnPEigen        cv::solvePnP(..., CV_ITERATIVE)   — one tag's four corners -> T_cam<-tag
//   solvePnPRansacEigen  cv::solvePnPRansac(...)           — corners of all reconstructed tags in an image -> camera pose
// OpenCV (C++) is not part of this build.  The same estimate is computed in-repo: undistort the observations,
// closed-form start (homography for coplanar points, DLT otherwise), then Levenberg-Marquardt on the pixel
// reprojection error of the full pinhole + radial-tangential model — the criterion CV_ITERATIVE minimises.
// The robust variant draws its hypotheses from the natural minimal sets of this problem — the four corners of each
// tag — instead of cv::solvePnPRansac's random 4-point samples (deterministic, no RNG), keeps the hypothesis with
// the most corners inside the same 8 px threshold and refits on those inliers.
// Host-side set-up code: it produces the start values the GPU bundle adjustment refines; it is not on the hot path.
#ifndef VISUAL_MARKER_MAPPING_PNP_H_
#define VISUAL_MARKER_MAPPING_PNP_H_
#include "visual_marker_mapping/CameraModel.h"
#include "visual_marker_mapping/EigenLite.h"
#include <vector>

namespace visual_marker_mapping
{
struct PnPResult
{
    Matrix3d R;          // object -> camera
    Vector3d t;
    double rmsError = 0; // pixel RMS of the reprojection error at the solution
    int iterations = 0;
    bool ok = false;
};

// normalised image coordinates (x, y) of a pixel: inverse of the distortion model by fixed-point iteration
Vector2d undistortPixel(const CameraModel& cam, const Vector2d& uv);

// pixel projection of a camera-frame point with the standard OpenCV model (the model of the cost functor,
// TagReconstructionCostFunction.h:125-152)
Vector2d projectPinhole(const CameraModel& cam, const Vector3d& pointInCamera);

// replaces solvePnPEigen (src/EigenCVConversions.cpp:38-67).  >= 4 correspondences (4 must be coplanar).
PnPResult solvePnP(const std::vector<Vector3d>& objectPoints, const std::vector<Vector2d>& observations, const CameraModel& cam);

// replaces solvePnPRansacEigen (src/EigenCVConversions.cpp:69-106).  Correspondences come in groups of four (the
// corners of one tag, as computeRelativeCameraPoseFromImg assembles them, src/TagReconstructor.cpp:286-301).
PnPResult solvePnPRobust(const std::vector<Vector3d>& objectPoints, const std::vector<Vector2d>& observations, const CameraModel& cam,
                         double inlierThresholdPx = 8.0, std::vector<unsigned char>* inliers = nullptr);

// rotation matrix -> quaternion (w,x,y,z), unit norm, w >= 0 not enforced (Eigen::Quaterniond(R) semantics)
Vector4d rotationMatrixToQuaternion(const Matrix3d& R);
}
#endif
