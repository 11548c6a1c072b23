// CameraUtilities.cpp — camera_intrinsics.json (src/CameraUtilities.cpp:45-67; README.md:130-151) and
// CameraModel::projectPoint (src/CameraModel.cpp:6-26).
#include "visual_marker_mapping/CameraUtilities.h"
#include "CameraModelJson.h"

#include <stdexcept>

namespace visual_marker_mapping
{
CameraModel cameraModelFromJson(const vmm_json::Node& n)
{
    CameraModel m;
    m.fx = n.get_double("fx");
    m.fy = n.get_double("fy");
    m.cx = n.get_double("cx");
    m.cy = n.get_double("cy");
    const auto& d = n.get_child("distortion_coefficients").children;
    // PropertyTreeUtilities.h:107-124: exactly 5 values
    if (d.size() > 5) throw std::runtime_error("Too many parameters in vector. Expected 5!");
    if (d.size() < 5) throw std::runtime_error("Not enough many parameters in vector. Expected 5, got " + std::to_string(d.size()) + "!");
    for (int i = 0; i < 5; ++i) m.distortionCoefficients[i] = d[i].second.value_double();
    m.horizontalResolution = n.get_int("horizontal_resolution");
    m.verticalResolution = n.get_int("vertical_resolution");
    return m;
}

vmm_json::Node cameraModelToJson(const CameraModel& m)
{
    vmm_json::Node n, d;
    n.put("fx", m.fx);
    n.put("fy", m.fy);
    n.put("cx", m.cx);
    n.put("cy", m.cy);
    for (int i = 0; i < 5; ++i) d.push_back(vmm_json::Node::leaf(m.distortionCoefficients[i]));
    n.add_child("distortion_coefficients", d);
    n.put("vertical_resolution", m.verticalResolution);
    n.put("horizontal_resolution", m.horizontalResolution);
    return n;
}

CameraModel readCameraModel(const std::string& filename)
{
    return cameraModelFromJson(vmm_json::read_json(filename));
}

Vector2d CameraModel::projectPoint(const Vector3d& point3D) const
{
    const double k1 = distortionCoefficients[0], k2 = distortionCoefficients[1], p1 = distortionCoefficients[2],
                 p2 = distortionCoefficients[3], k3 = distortionCoefficients[4];
    double x = point3D.x() / point3D.z();
    double y = point3D.y() / point3D.z();
    const double r2 = x * x + y * y;
    // the reference overwrites x before computing y (src/CameraModel.cpp:20-23); kept, the culling thresholds depend on it
    x = x * (1 + r2 * (k1 + r2 * (k2 + r2 * k3))) + 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
    y = y * (1 + r2 * (k1 + r2 * (k2 + r2 * k3))) + 2 * p2 * x * y + p1 * (r2 + 2 * y * y);
    return Vector2d(fx * x + cx, fy * y + cy);
}
}
