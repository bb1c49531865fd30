// scene_xml.cpp -- collision geometry from an OpenRAVE environment XML.
//
// The reference never parses its scene files itself: OpenRAVE loads them (MCSimulation.py:123-124) and the
// plugin asks OpenRAVE for collisions (mcsimplugin/MCSimulator.h:269-285).  This path replaces that query with
// SAT on z-extruded oriented rectangles, so the rectangles are extracted here, once, when the module is created.
// Semantics, all visible in pr2custom.env.xml:14-249 and pr2test2.env.xml:12-122:
//   <Extents> half-extents; <Translation> the geom centre; <RotationAxis>ax ay az deg</RotationAxis> a rotation
//   about the geom's own centre (only about +-z is representable as a z-prism); <offsetfrom>Body</offsetfrom> the
//   frame of a sibling body; KinBody-level <Translation>/<RotationAxis> move the whole body; file="..." pulls
//   geometry from another file, resolved next to the env file and then on the fixture path (ikeatable.kinbody.xml
//   and the PR2 model are not in the reference repository); tag case varies (<translation>, pr2test2.env.xml:121).
//   A <Robot file> that resolves to a COLLADA model (.zae / .dae, what the reference's scenes name) is reduced to convex
//   footprints per z-band by collada_hulls.cpp; when only the box fixture is there, that is read instead.
// Beyond the reference's files (robot hull sets that are not boxes): <Geom type="prism"> with
//   <Vertices>x0 y0 x1 y1 ...</Vertices> (3..8 vertices of a convex polygon, either orientation) and <Height>zlo zhi
//   </Height> in the geom's frame; accepted inside <Robot> only (obstacles stay boxes).
// Output rows (8 doubles): cx, cy, cos(yaw), sin(yaw), hx, hy, zlo, zhi -- obstacles in the world frame, robot hulls
// in the robot frame with absolute z (the robot's z offset from the env file applied).
#include "scene_xml.h"

#include <cmath>
#include <fstream>
#include <map>
#include <sstream>

#include "collada_hulls.h"
#include "xml_lite.h"

namespace mcsim {
namespace {

constexpr double kPi = 3.14159265358979323846;

struct Frame {  // planar rigid transform + z offset
    double x = 0, y = 0, z = 0, yaw = 0;
    Frame then(const Frame& o) const {  // this * o
        const double c = std::cos(yaw), s = std::sin(yaw);
        Frame r;
        r.x = x + c * o.x - s * o.y;
        r.y = y + s * o.x + c * o.y;
        r.z = z + o.z;
        r.yaw = yaw + o.yaw;
        return r;
    }
};

std::vector<double> numbers(const std::string& text, const char* what) {
    std::istringstream is(text);
    std::vector<double> v;
    double d;
    while (is >> d) v.push_back(d);
    if (v.empty()) throw std::runtime_error(std::string("no numbers in <") + what + ">");
    return v;
}

Frame local_frame(const xml_lite::Node& el) {
    Frame f;
    for (const auto* t : el.all("translation")) {
        const auto v = numbers(t->text, "Translation");
        if (v.size() < 3) throw std::runtime_error("<Translation> needs 3 numbers");
        f.x = v[0]; f.y = v[1]; f.z = v[2];
    }
    for (const auto* r : el.all("rotationaxis")) {
        const auto v = numbers(r->text, "RotationAxis");
        if (v.size() < 4) throw std::runtime_error("<RotationAxis> needs 4 numbers");
        const double n = std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        if (n == 0.0 || v[3] == 0.0) continue;
        if (std::fabs(v[0]) > 1e-12 * n || std::fabs(v[1]) > 1e-12 * n)
            throw std::runtime_error("<RotationAxis>" + r->text + "</RotationAxis> is not about z: not a z-prism");
        f.yaw = v[3] * kPi / 180.0 * (v[2] > 0 ? 1.0 : -1.0);
    }
    return f;
}

std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::ostringstream os;
    os << f.rdbuf();
    return os.str();
}

bool readable(const std::string& path) { return bool(std::ifstream(path)); }

bool ends_with(const std::string& s, const std::string& suf) {
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}

std::string resolve(const std::string& ref, const std::string& base_dir, const std::vector<std::string>& fixture_dirs) {
    std::vector<std::string> cands{base_dir + "/" + ref};
    const size_t slash = ref.find_last_of('/');
    const std::string stem = slash == std::string::npos ? ref : ref.substr(slash + 1);
    const size_t dot = stem.find_last_of('.');
    const std::string bare = dot == std::string::npos ? stem : stem.substr(0, dot);
    for (const auto& d : fixture_dirs) {
        cands.push_back(d + "/" + stem);
        cands.push_back(d + "/" + bare + ".fixture.xml");
    }
    std::string tried;
    for (const auto& c : cands) {
        if (readable(c)) return c;   // a COLLADA model (.zae / .dae) that is really there wins over the box fixture
        tried += (tried.empty() ? "" : ", ") + c;
    }
    throw std::runtime_error("cannot resolve '" + ref + "' (tried " + tried + ")");
}

struct Prism { Frame f; double hx, hy, hz; };
struct PolyPrism { Frame f; std::vector<double> xy; double zlo, zhi; };

// every box geom of a <KinBody>/<Robot> element, in the frame `parent`
void collect(const xml_lite::Node& kin, const Frame& parent, std::vector<Prism>& out, std::vector<PolyPrism>* polys = nullptr) {
    std::map<std::string, Frame> frames;
    for (const auto* body : kin.all("body")) {
        Frame f = local_frame(*body);
        if (const auto* off = body->first("offsetfrom")) {
            std::string name = off->text;
            name.erase(0, name.find_first_not_of(" \t\r\n"));
            name.erase(name.find_last_not_of(" \t\r\n") + 1);
            const auto it = frames.find(name);
            if (it == frames.end()) throw std::runtime_error("<offsetfrom>" + name + "</offsetfrom>: no such body before it");
            f = it->second.then(f);
        }
        if (const auto* nm = body->attr("name")) frames[*nm] = f;
        for (const auto* geom : body->all("geom")) {
            const auto* type = geom->attr("type");
            if (type && xml_lite::lower(*type) == "prism" && polys) {
                const auto* vt = geom->first("vertices");
                const auto* ht = geom->first("height");
                if (!vt || !ht) throw std::runtime_error("prism geom needs <Vertices> and <Height>");
                const auto v = numbers(vt->text, "Vertices"), hz = numbers(ht->text, "Height");
                if (v.size() < 6 || v.size() > 16 || v.size() % 2) throw std::runtime_error("<Vertices> needs 3..8 x y pairs");
                if (hz.size() < 2 || !(hz[0] < hz[1])) throw std::runtime_error("<Height> needs zlo < zhi");
                polys->push_back(PolyPrism{parent.then(f).then(local_frame(*geom)), v, hz[0], hz[1]});
                continue;
            }
            if (!type || xml_lite::lower(*type) != "box")
                throw std::runtime_error("only box geoms (and prism geoms in a <Robot>) are supported, got type='" + (type ? *type : std::string()) + "'");
            const auto* ext = geom->first("extents");
            if (!ext) throw std::runtime_error("box geom without <Extents>");
            const auto e = numbers(ext->text, "Extents");
            if (e.size() < 3) throw std::runtime_error("<Extents> needs 3 numbers");
            out.push_back(Prism{parent.then(f).then(local_frame(*geom)), e[0], e[1], e[2]});
        }
    }
}

// vertices into the parent frame, counter-clockwise
void push_poly(std::vector<double>& rows, const PolyPrism& p) {
    const size_t n = p.xy.size() / 2;
    std::vector<double> w(2 * n);
    const double c = std::cos(p.f.yaw), s = std::sin(p.f.yaw);
    double area2 = 0;
    for (size_t i = 0; i < n; ++i) {
        w[2 * i] = p.f.x + c * p.xy[2 * i] - s * p.xy[2 * i + 1];
        w[2 * i + 1] = p.f.y + s * p.xy[2 * i] + c * p.xy[2 * i + 1];
    }
    for (size_t i = 0; i < n; ++i) { const size_t j = (i + 1) % n; area2 += w[2 * i] * w[2 * j + 1] - w[2 * j] * w[2 * i + 1]; }
    double r[19] = {double(n), p.f.z + p.zlo, p.f.z + p.zhi};
    for (size_t i = 0; i < n; ++i) {
        const size_t k = area2 >= 0 ? i : n - 1 - i;   // clockwise input: reversed
        r[3 + 2 * i] = w[2 * k]; r[4 + 2 * i] = w[2 * k + 1];
    }
    rows.insert(rows.end(), r, r + 19);
}

void push_row(std::vector<double>& rows, const Prism& p) {
    const double r[8] = {p.f.x, p.f.y, std::cos(p.f.yaw), std::sin(p.f.yaw), p.hx, p.hy, p.f.z - p.hz, p.f.z + p.hz};
    rows.insert(rows.end(), r, r + 8);
}

}  // namespace

Scene load_scene_xml(const std::string& env_path, const std::vector<std::string>& fixture_dirs) {
    const size_t slash = env_path.find_last_of('/');
    const std::string base_dir = slash == std::string::npos ? "." : env_path.substr(0, slash);
    const std::string text = read_file(env_path);
    const auto root = xml_lite::parse(text);
    Scene sc;
    for (const auto* kin : root->all("kinbody")) {
        const Frame top = local_frame(*kin);
        std::vector<Prism> prisms;
        if (const auto* file = kin->attr("file")) {
            const std::string sub = read_file(resolve(*file, base_dir, fixture_dirs));
            collect(*xml_lite::parse(sub), top, prisms);
        } else {
            collect(*kin, top, prisms);
        }
        for (const auto& p : prisms) push_row(sc.boxes, p);
    }
    const auto robots = root->all("robot");
    if (!robots.empty()) {
        const auto* rob = robots.front();
        const Frame top = local_frame(*rob);
        sc.robot_pose[0] = top.x; sc.robot_pose[1] = top.y; sc.robot_pose[2] = top.yaw;
        sc.has_robot = true;
        Frame zonly;
        zonly.z = top.z;  // hulls stay in the robot frame; only the robot's height is applied
        std::vector<Prism> prisms;
        std::vector<PolyPrism> polys;
        if (const auto* file = rob->attr("file")) {
            const std::string path = resolve(*file, base_dir, fixture_dirs);
            if (ends_with(path, ".zae") || ends_with(path, ".dae")) {
                // the COLLADA model itself: one convex footprint per z-band (collada_hulls.h).  zbands="z0 z1 .."
                // on the <Robot> element (not OpenRAVE's: this build's) gives the band boundaries in the model frame.
                // hulls="boxes" replaces every footprint by its minimum-area bounding rectangle (box hulls: the packed kernels).
                ColladaOptions opt;
                if (const auto* zb = rob->attr("zbands")) opt.z_edges = numbers(*zb, "Robot zbands");
                if (const auto* hk = rob->attr("hulls")) {
                    if (xml_lite::lower(*hk) == "boxes") { opt.as_boxes = true; opt.max_bands = 8; }
                    else if (xml_lite::lower(*hk) != "polygons") throw std::runtime_error("<Robot hulls=\"" + *hk + "\">: boxes | polygons expected");
                }
                std::vector<double> rows = load_collada_hulls(path, opt);
                if (opt.as_boxes) {
                    for (size_t i = 0; i < rows.size(); i += 8) { rows[i + 6] += top.z; rows[i + 7] += top.z; }
                    sc.hulls.insert(sc.hulls.end(), rows.begin(), rows.end());
                } else {
                    for (size_t i = 0; i < rows.size(); i += 19) { rows[i + 1] += top.z; rows[i + 2] += top.z; }
                    sc.poly_hulls.insert(sc.poly_hulls.end(), rows.begin(), rows.end());
                }
            } else {
                collect(*xml_lite::parse(read_file(path)), zonly, prisms, &polys);
            }
        } else {
            collect(*rob, zonly, prisms, &polys);
        }
        for (const auto& p : prisms) push_row(sc.hulls, p);
        for (const auto& p : polys) push_poly(sc.poly_hulls, p);
    }
    return sc;
}

}  // namespace mcsim
