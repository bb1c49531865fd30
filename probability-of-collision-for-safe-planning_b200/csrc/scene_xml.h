// scene_xml.h -- obstacle / robot prisms extracted from an OpenRAVE environment XML (see scene_xml.cpp).
#ifndef MCSIM_SCENE_XML_H
#define MCSIM_SCENE_XML_H

#include <stdexcept>
#include <string>
#include <vector>

namespace mcsim {

struct Scene {
    std::vector<double> boxes;  // n x 8: cx, cy, cos(yaw), sin(yaw), hx, hy, zlo, zhi   (world frame)
    std::vector<double> hulls;  // m x 8: same layout in the robot frame, z absolute
    // robot hulls that are not boxes: <Geom type="prism"> = a convex polygon footprint extruded in z (robot frame, z
    // absolute); k x 19: n, zlo, zhi, x0, y0, ..., x7, y7 (counter-clockwise, at most 8 vertices)
    std::vector<double> poly_hulls;
    int n_poly_hulls() const { return int(poly_hulls.size() / 19); }
    double robot_pose[3] = {0, 0, 0};
    bool has_robot = false;
    int n_boxes() const { return int(boxes.size() / 8); }
    int n_hulls() const { return int(hulls.size() / 8); }
};

// Throws std::runtime_error with a message naming the file / tag on any problem.
Scene load_scene_xml(const std::string& env_path, const std::vector<std::string>& fixture_dirs);

}  // namespace mcsim
#endif
