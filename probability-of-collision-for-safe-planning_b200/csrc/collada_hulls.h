// collada_hulls.h -- robot hull set from a COLLADA model (.dae) or archive (.zae), SURVEY.md 8f-3.
//
// The reference's scenes name their robot as a COLLADA archive (pr2test2.env.xml:120, pr2custom.env.xml:247:
// <Robot file="robots/pr2-beta-static.zae">); OpenRAVE loads the meshes and its collision checker tests them
// (mcsimplugin/MCSimulator.h:269-285).  This path tests 2.5-D prisms, so the model is reduced to what that needs: the
// triangles of every instantiated geometry are brought into the model's root frame (visual-scene transforms, unit and
// up-axis of <asset>), cut into horizontal bands, and each band's footprint -- the convex hull of the xy-projection of the
// part of every triangle inside the band -- becomes one convex-polygon hull (mcsim_cuda_set_scene_ex), conservatively
// simplified to at most 8 vertices.  A static model only: joints keep the pose the visual scene gives them.
#ifndef MCSIM_COLLADA_HULLS_H
#define MCSIM_COLLADA_HULLS_H

#include <string>
#include <vector>

namespace mcsim {

struct ColladaOptions {
    std::vector<double> z_edges;   // band boundaries z0 < z1 < ... (model frame, metres); empty: max_bands equal bands over the model's height
    int max_bands = 4;             // = MCSIM_MAX_POLY_HULLS
    int max_vertices = 8;          // = MCSIM_POLY_MAX_VERTICES
    bool as_boxes = false;         // rows of 8 doubles instead: each band's footprint replaced by its minimum-area bounding rectangle
                                   // (cx, cy, cos(yaw), sin(yaw), hx, hy, zlo, zhi) -- looser than the polygon, but box hulls run on the
                                   // packed kernels (3-4 x the rollout rate of polygon hulls)
};

// Rows of 19 doubles, one per non-empty band: n, zlo, zhi, x0, y0, ..., x7, y7 (counter-clockwise, model frame); with
// opt.as_boxes rows of 8 doubles (the box hull layout of mcsim_cuda_set_scene).
// Throws std::runtime_error naming the file and the construct on any problem.
std::vector<double> load_collada_hulls(const std::string& path, const ColladaOptions& opt);

// The pieces, exposed for tests: the .dae text inside a .zae (zip: stored or deflated entries; the entry named by
// manifest.xml's <dae_root>, else the first *.dae), and the hull rows of a .dae text.
std::string read_zae_document(const std::string& path);
std::vector<double> collada_hulls_from_text(const std::string& dae_text, const ColladaOptions& opt, const std::string& what);

}  // namespace mcsim
#endif
