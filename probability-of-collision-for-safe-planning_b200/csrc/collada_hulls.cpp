// collada_hulls.cpp -- see collada_hulls.h.  COLLADA 1.4 / 1.5 subset: <asset> unit + up_axis, <library_geometries>
// meshes (<triangles>, <polylist>, <polygons>; other primitives contribute their vertices), <library_nodes>,
// <library_visual_scenes> node trees with <matrix> / <translate> / <rotate> / <scale>, <instance_geometry>,
// <instance_node>.  Kinematics, physics, materials and controllers are not read.
#include "collada_hulls.h"

#include <zlib.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>

#include "xml_lite.h"

namespace mcsim {
namespace {

[[noreturn]] void fail(const std::string& what, const std::string& msg) { throw std::runtime_error(what + ": " + msg); }

std::string slurp(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

bool ends_with_ci(const std::string& s, const char* suf) {
    const size_t n = std::strlen(suf);
    if (s.size() < n) return false;
    for (size_t i = 0; i < n; ++i)
        if (std::tolower(static_cast<unsigned char>(s[s.size() - n + i])) != suf[i]) return false;
    return true;
}

// ---- zip (PKWARE APPNOTE 4.3: end of central directory, central file headers, local file headers) -----------------
uint32_t rd32(const std::string& b, size_t o) {
    return uint32_t(uint8_t(b[o])) | uint32_t(uint8_t(b[o + 1])) << 8 | uint32_t(uint8_t(b[o + 2])) << 16 | uint32_t(uint8_t(b[o + 3])) << 24;
}
uint16_t rd16(const std::string& b, size_t o) { return uint16_t(uint8_t(b[o]) | uint8_t(b[o + 1]) << 8); }

struct ZipEntry { std::string name; uint16_t method; uint32_t csize, usize, local; };

std::vector<ZipEntry> zip_entries(const std::string& b, const std::string& what) {
    if (b.size() < 22) fail(what, "not a zip archive (too short)");
    size_t eocd = std::string::npos;
    for (size_t back = 22; back <= b.size() && back <= 22 + 65535; ++back)
        if (rd32(b, b.size() - back) == 0x06054b50u) { eocd = b.size() - back; break; }
    if (eocd == std::string::npos) fail(what, "not a zip archive (no end-of-central-directory record)");
    const unsigned n = rd16(b, eocd + 10);
    size_t o = rd32(b, eocd + 16);
    std::vector<ZipEntry> out;
    for (unsigned i = 0; i < n; ++i) {
        if (o + 46 > b.size() || rd32(b, o) != 0x02014b50u) fail(what, "corrupt zip central directory");
        ZipEntry e;
        e.method = rd16(b, o + 10);
        e.csize = rd32(b, o + 20);
        e.usize = rd32(b, o + 24);
        const unsigned nl = rd16(b, o + 28), xl = rd16(b, o + 30), cl = rd16(b, o + 32);
        e.local = rd32(b, o + 42);
        if (o + 46 + nl > b.size()) fail(what, "corrupt zip central directory");
        e.name = b.substr(o + 46, nl);
        out.push_back(e);
        o += 46 + nl + xl + cl;
    }
    return out;
}

std::string zip_extract(const std::string& b, const ZipEntry& e, const std::string& what) {
    if (size_t(e.local) + 30 > b.size() || rd32(b, e.local) != 0x04034b50u) fail(what, "corrupt zip local header of " + e.name);
    const size_t data = size_t(e.local) + 30 + rd16(b, e.local + 26) + rd16(b, e.local + 28);
    if (data + e.csize > b.size()) fail(what, "zip entry " + e.name + " runs past the end of the archive");
    if (e.method == 0) return b.substr(data, e.csize);
    if (e.method != 8) fail(what, "zip entry " + e.name + " uses compression method " + std::to_string(e.method) + " (stored and deflate are read)");
    std::string out(e.usize, '\0');
    z_stream zs;
    std::memset(&zs, 0, sizeof(zs));
    if (inflateInit2(&zs, -MAX_WBITS) != Z_OK) fail(what, "zlib inflateInit2 failed");
    zs.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(b.data() + data));
    zs.avail_in = e.csize;
    zs.next_out = reinterpret_cast<Bytef*>(&out[0]);
    zs.avail_out = e.usize;
    const int rc = inflate(&zs, Z_FINISH);
    inflateEnd(&zs);
    if (rc != Z_STREAM_END || zs.total_out != e.usize) fail(what, "zip entry " + e.name + " does not inflate to its stated size");
    return out;
}

// ---- small linear algebra ----------------------------------------------------------------------------------------------
struct M4 {
    double m[16];
    static M4 identity() { M4 r; std::memset(r.m, 0, sizeof(r.m)); r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1; return r; }
    M4 operator*(const M4& o) const {
        M4 r;
        for (int i = 0; i < 4; ++i)
            for (int j = 0; j < 4; ++j) {
                double s = 0;
                for (int k = 0; k < 4; ++k) s += m[4 * i + k] * o.m[4 * k + j];
                r.m[4 * i + j] = s;
            }
        return r;
    }
    void apply(const double* p, double* q) const {
        for (int i = 0; i < 3; ++i) q[i] = m[4 * i] * p[0] + m[4 * i + 1] * p[1] + m[4 * i + 2] * p[2] + m[4 * i + 3];
    }
};

std::vector<double> doubles(const std::string& text) {
    std::vector<double> v;
    const char* p = text.c_str();
    char* e = nullptr;
    for (;;) {
        const double d = std::strtod(p, &e);
        if (e == p) break;
        v.push_back(d);
        p = e;
    }
    return v;
}
std::vector<long> longs(const std::string& text) {
    std::vector<long> v;
    const char* p = text.c_str();
    char* e = nullptr;
    for (;;) {
        const long d = std::strtol(p, &e, 10);
        if (e == p) break;
        v.push_back(d);
        p = e;
    }
    return v;
}
std::string frag(const std::string& url) { return !url.empty() && url[0] == '#' ? url.substr(1) : url; }

struct Tri { double v[9]; };

// ---- the document ------------------------------------------------------------------------------------------------------
struct Doc {
    std::string what;
    std::map<std::string, std::vector<Tri>> geometry;          // id -> triangles in the geometry's own frame
    std::map<std::string, const xml_lite::Node*> nodes;        // every <node> with an id
    std::vector<Tri> out;                                      // instantiated triangles, root frame (document units)
    int depth = 0;

    void index_nodes(const xml_lite::Node& n) {
        for (const auto& c : n.children) {
            if (c->ltag == "node") {
                if (const auto* id = c->attr("id")) nodes[*id] = c.get();
                index_nodes(*c);
            }
        }
    }

    void read_geometry(const xml_lite::Node& g) {
        const auto* id = g.attr("id");
        const auto* mesh = g.first("mesh");
        if (!id || !mesh) return;   // <convex_mesh>, <spline>: not collision geometry this path reads
        std::map<std::string, std::pair<std::vector<double>, int>> sources;   // id -> (floats, stride)
        for (const auto* s : mesh->all("source")) {
            const auto* sid = s->attr("id");
            const auto* fa = s->first("float_array");
            if (!sid || !fa) continue;
            int stride = 3;
            if (const auto* tc = s->first("technique_common"))
                if (const auto* ac = tc->first("accessor"))
                    if (const auto* st = ac->attr("stride")) stride = std::max(1, atoi(st->c_str()));
            sources[*sid] = std::make_pair(doubles(fa->text), stride);
        }
        std::map<std::string, std::string> vertices;   // <vertices id> -> POSITION source id
        for (const auto* v : mesh->all("vertices")) {
            const auto* vid = v->attr("id");
            if (!vid) continue;
            for (const auto* in : v->all("input")) {
                const auto* sem = in->attr("semantic");
                const auto* src = in->attr("source");
                if (sem && src && *sem == "POSITION") vertices[*vid] = frag(*src);
            }
        }
        std::vector<Tri>& tris = geometry[*id];
        auto position = [&](const std::pair<std::vector<double>, int>& src, long idx, double* p) {
            const size_t o = size_t(idx) * src.second;
            if (idx < 0 || o + 3 > src.first.size()) fail(what, "geometry " + *id + ": vertex index " + std::to_string(idx) + " out of range");
            p[0] = src.first[o]; p[1] = src.first[o + 1]; p[2] = src.first[o + 2];
        };
        for (const auto& prim : mesh->children) {
            const std::string& k = prim->ltag;
            if (k != "triangles" && k != "polylist" && k != "polygons" && k != "tristrips" && k != "trifans" && k != "lines" && k != "linestrips") continue;
            int stride = 0, voff = -1;
            std::string vsrc;
            for (const auto* in : prim->all("input")) {
                const auto* sem = in->attr("semantic");
                const auto* off = in->attr("offset");
                const auto* src = in->attr("source");
                const int o = off ? atoi(off->c_str()) : 0;
                stride = std::max(stride, o + 1);
                if (sem && src && *sem == "VERTEX") { voff = o; vsrc = frag(*src); }
            }
            if (voff < 0) continue;
            const auto vit = vertices.find(vsrc);
            if (vit == vertices.end()) fail(what, "geometry " + *id + ": <" + prim->tag + "> refers to unknown <vertices> " + vsrc);
            const auto sit = sources.find(vit->second);
            if (sit == sources.end()) fail(what, "geometry " + *id + ": no <source> " + vit->second);
            const auto& src = sit->second;
            auto emit = [&](long a, long b, long c) {
                Tri t;
                position(src, a, t.v); position(src, b, t.v + 3); position(src, c, t.v + 6);
                tris.push_back(t);
            };
            std::vector<long> vcount;
            if (k == "polylist")
                if (const auto* vc = prim->first("vcount")) vcount = longs(vc->text);
            size_t poly = 0;
            for (const auto* pn : prim->all("p")) {
                const std::vector<long> p = longs(pn->text);
                const size_t nv = p.size() / size_t(stride);
                auto vi = [&](size_t j) { return p[j * stride + voff]; };
                if (k == "triangles") {
                    for (size_t j = 0; j + 2 < nv; j += 3) emit(vi(j), vi(j + 1), vi(j + 2));
                } else if (k == "polylist") {
                    size_t j = 0;
                    for (; poly < vcount.size() && j + size_t(vcount[poly]) <= nv; ++poly) {
                        const size_t c = size_t(vcount[poly]);
                        for (size_t q = 1; q + 1 < c; ++q) emit(vi(j), vi(j + q), vi(j + q + 1));   // fan
                        if (c < 3) for (size_t q = 0; q < c; ++q) emit(vi(j + q), vi(j + q), vi(j + q));
                        j += c;
                    }
                } else if (k == "polygons" || k == "trifans") {
                    for (size_t q = 1; q + 1 < nv; ++q) emit(vi(0), vi(q), vi(q + 1));
                } else if (k == "tristrips") {
                    for (size_t q = 0; q + 2 < nv; ++q) emit(vi(q), vi(q + 1), vi(q + 2));
                } else {   // lines: their vertices still bound the shape
                    for (size_t q = 0; q < nv; ++q) emit(vi(q), vi(q), vi(q));
                }
            }
        }
    }

    static M4 transform_of(const xml_lite::Node& el, const std::string& what) {
        M4 t = M4::identity();
        const std::vector<double> v = doubles(el.text);
        if (el.ltag == "matrix") {
            if (v.size() < 16) fail(what, "<matrix> needs 16 numbers");
            for (int i = 0; i < 16; ++i) t.m[i] = v[i];   // row-major, column vectors (COLLADA 1.4.1, 5-51)
        } else if (el.ltag == "translate") {
            if (v.size() < 3) fail(what, "<translate> needs 3 numbers");
            t.m[3] = v[0]; t.m[7] = v[1]; t.m[11] = v[2];
        } else if (el.ltag == "scale") {
            if (v.size() < 3) fail(what, "<scale> needs 3 numbers");
            t.m[0] = v[0]; t.m[5] = v[1]; t.m[10] = v[2];
        } else if (el.ltag == "rotate") {
            if (v.size() < 4) fail(what, "<rotate> needs axis + angle in degrees");
            double x = v[0], y = v[1], z = v[2];
            const double n = std::sqrt(x * x + y * y + z * z);
            if (n > 0) {
                x /= n; y /= n; z /= n;
                const double a = v[3] * 3.14159265358979323846 / 180.0, c = std::cos(a), s = std::sin(a), k = 1 - c;
                const double r[9] = {c + x * x * k, x * y * k - z * s, x * z * k + y * s,
                                     y * x * k + z * s, c + y * y * k, y * z * k - x * s,
                                     z * x * k - y * s, z * y * k + x * s, c + z * z * k};
                for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) t.m[4 * i + j] = r[3 * i + j];
            }
        }
        return t;
    }

    void walk(const xml_lite::Node& node, M4 m) {
        if (++depth > 256) fail(what, "node hierarchy deeper than 256 (a cycle of <instance_node>?)");
        for (const auto& c : node.children) {
            const std::string& k = c->ltag;
            if (k == "matrix" || k == "translate" || k == "rotate" || k == "scale") {
                m = m * transform_of(*c, what);   // transforms compose in document order (COLLADA 1.4.1, 5-86)
            } else if (k == "instance_geometry") {
                const auto* url = c->attr("url");
                if (!url) continue;
                const auto it = geometry.find(frag(*url));
                if (it == geometry.end()) continue;   // a geometry without a <mesh>
                for (const Tri& t : it->second) {
                    Tri w;
                    for (int q = 0; q < 3; ++q) m.apply(t.v + 3 * q, w.v + 3 * q);
                    out.push_back(w);
                }
            } else if (k == "instance_node") {
                const auto* url = c->attr("url");
                if (!url) continue;
                const auto it = nodes.find(frag(*url));
                if (it == nodes.end()) fail(what, "<instance_node url=\"" + *url + "\">: no such node");
                walk(*it->second, m);
            } else if (k == "node") {
                walk(*c, m);
            }
        }
        --depth;
    }
};

// ---- footprints --------------------------------------------------------------------------------------------------------
struct P2 { double x, y; };
double cross(const P2& o, const P2& a, const P2& b) { return (a.x - o.x) * (b.y - o.y) - (a.y - o.y) * (b.x - o.x); }

std::vector<P2> convex_hull(std::vector<P2> p) {   // Andrew's monotone chain; counter-clockwise, no collinear vertices
    std::sort(p.begin(), p.end(), [](const P2& a, const P2& b) { return a.x < b.x || (a.x == b.x && a.y < b.y); });
    p.erase(std::unique(p.begin(), p.end(), [](const P2& a, const P2& b) { return a.x == b.x && a.y == b.y; }), p.end());
    if (p.size() < 3) return p;
    std::vector<P2> h(2 * p.size());
    size_t k = 0;
    for (size_t i = 0; i < p.size(); ++i) {
        while (k >= 2 && cross(h[k - 2], h[k - 1], p[i]) <= 0) --k;
        h[k++] = p[i];
    }
    for (size_t i = p.size() - 1, t = k + 1; i-- > 0;) {
        while (k >= t && cross(h[k - 2], h[k - 1], p[i]) <= 0) --k;
        h[k++] = p[i];
    }
    h.resize(k - 1);
    // points cut out of the same edge by different triangles are collinear only up to rounding: drop the vertices whose
    // turn is below 1e-12 of the product of the adjacent edge lengths
    for (bool again = h.size() > 3; again;) {
        again = false;
        for (size_t i = 0; i < h.size() && h.size() > 3; ++i) {
            const P2 &a = h[(i + h.size() - 1) % h.size()], &b = h[i], &c = h[(i + 1) % h.size()];
            const double ab = std::hypot(b.x - a.x, b.y - a.y), bc = std::hypot(c.x - b.x, c.y - b.y);
            if (std::fabs(cross(a, b, c)) <= 1e-12 * ab * bc || ab == 0 || bc == 0) { h.erase(h.begin() + long(i)); again = true; break; }
        }
    }
    return h;
}

// A convex superset with one vertex less: drop the edge (v_i, v_{i+1}) whose two neighbours, extended, meet closest to it
// (the added triangle v_i, p, v_{i+1} has the least area).  For n >= 5 some edge always has converging neighbours.
bool drop_cheapest_edge(std::vector<P2>& h) {
    const size_t n = h.size();
    double best = 1e300;
    size_t bi = n;
    P2 bp{0, 0};
    for (size_t i = 0; i < n; ++i) {
        const P2 &a = h[(i + n - 1) % n], &b = h[i], &c = h[(i + 1) % n], &d = h[(i + 2) % n];
        const double rx = b.x - a.x, ry = b.y - a.y, sx = c.x - d.x, sy = c.y - d.y;   // rays a->b and d->c
        const double den = rx * sy - ry * sx;
        if (std::fabs(den) < 1e-300) continue;
        const double t = ((d.x - a.x) * sy - (d.y - a.y) * sx) / den, u = ((d.x - a.x) * ry - (d.y - a.y) * rx) / den;
        if (!(t > 1.0) || !(u > 1.0)) continue;   // the extensions must meet beyond b and beyond c
        const P2 p{a.x + t * rx, a.y + t * ry};
        const double area = 0.5 * std::fabs(cross(b, p, c));
        if (area < best) { best = area; bi = i; bp = p; }
    }
    if (bi == n) return false;
    h[bi] = bp;
    h.erase(h.begin() + long((bi + 1) % n));
    return true;
}

// the part of a triangle with lo <= z <= hi (Sutherland-Hodgman against the two planes); returns its vertices
int clip_band(const Tri& t, double lo, double hi, double out[7][3]) {
    double a[7][3], b[7][3];
    int n = 3;
    for (int i = 0; i < 3; ++i) for (int k = 0; k < 3; ++k) a[i][k] = t.v[3 * i + k];
    for (int pass = 0; pass < 2; ++pass) {
        const double plane = pass == 0 ? lo : hi, sgn = pass == 0 ? 1.0 : -1.0;
        int m = 0;
        for (int i = 0; i < n; ++i) {
            const double* p = a[i];
            const double* q = a[(i + 1) % n];
            const double dp = sgn * (p[2] - plane), dq = sgn * (q[2] - plane);
            if (dp >= 0) { for (int k = 0; k < 3; ++k) b[m][k] = p[k]; ++m; }
            if ((dp > 0 && dq < 0) || (dp < 0 && dq > 0)) {
                const double s = dp / (dp - dq);
                for (int k = 0; k < 3; ++k) b[m][k] = p[k] + s * (q[k] - p[k]);
                b[m][2] = plane;
                ++m;
            }
        }
        n = m;
        for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) a[i][k] = b[i][k];
        if (n == 0) break;
    }
    for (int i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) out[i][k] = a[i][k];
    return n;
}

}  // namespace

std::string read_zae_document(const std::string& path) {
    const std::string b = slurp(path);
    const std::vector<ZipEntry> es = zip_entries(b, path);
    std::string want;
    for (const auto& e : es)
        if (e.name == "manifest.xml") {
            const auto root = xml_lite::parse(zip_extract(b, e, path));
            if (const auto* r = root->first("dae_root")) {
                want = r->text;
                want.erase(0, want.find_first_not_of(" \t\r\n./"));
                want.erase(want.find_last_not_of(" \t\r\n") + 1);
            }
        }
    const ZipEntry* pick = nullptr;
    for (const auto& e : es) if (!want.empty() && e.name == want) pick = &e;
    if (!pick) for (const auto& e : es) if (ends_with_ci(e.name, ".dae")) { pick = &e; break; }
    if (!pick) fail(path, "no .dae document in the archive");
    return zip_extract(b, *pick, path);
}

std::vector<double> collada_hulls_from_text(const std::string& text, const ColladaOptions& opt, const std::string& what) {
    const auto root = xml_lite::parse(text);
    if (root->ltag != "collada") fail(what, "root element is <" + root->tag + ">, not <COLLADA>");
    double meter = 1.0;
    std::string up = "Y_UP";   // the COLLADA default
    if (const auto* asset = root->first("asset")) {
        if (const auto* u = asset->first("unit")) if (const auto* m = u->attr("meter")) meter = atof(m->c_str());
        if (const auto* a = asset->first("up_axis")) {
            up = a->text;
            up.erase(0, up.find_first_not_of(" \t\r\n"));
            up.erase(up.find_last_not_of(" \t\r\n") + 1);
        }
    }
    if (!(meter > 0)) fail(what, "<unit meter> must be positive");
    if (up != "Z_UP" && up != "Y_UP") fail(what, "<up_axis>" + up + "</up_axis> is not supported (Z_UP, Y_UP)");
    Doc d;
    d.what = what;
    for (const auto* lib : root->all("library_geometries"))
        for (const auto* g : lib->all("geometry")) d.read_geometry(*g);
    for (const auto* lib : root->all("library_nodes")) d.index_nodes(*lib);
    const xml_lite::Node* vs = nullptr;
    bool vs_matched = false;
    std::string vs_id;
    if (const auto* sc = root->first("scene"))
        if (const auto* iv = sc->first("instance_visual_scene"))
            if (const auto* url = iv->attr("url")) vs_id = frag(*url);
    for (const auto* lib : root->all("library_visual_scenes")) {
        d.index_nodes(*lib);
        for (const auto* v : lib->all("visual_scene")) {
            d.index_nodes(*v);
            const auto* id = v->attr("id");
            const bool match = id && !vs_id.empty() && *id == vs_id;
            if ((match && !vs_matched) || !vs) { vs = v; vs_matched = match; }   // the one <scene> instantiates, else the first
        }
    }
    if (!vs) fail(what, "no <visual_scene>");
    d.walk(*vs, M4::identity());
    if (d.out.empty()) fail(what, "the visual scene instantiates no mesh geometry");
    // root frame, metres, z up
    for (Tri& t : d.out)
        for (int q = 0; q < 3; ++q) {
            double* p = t.v + 3 * q;
            const double x = p[0] * meter, y = p[1] * meter, z = p[2] * meter;
            if (up == "Y_UP") { p[0] = x; p[1] = -z; p[2] = y; } else { p[0] = x; p[1] = y; p[2] = z; }
        }
    double zmin = 1e300, zmax = -1e300;
    for (const Tri& t : d.out) for (int q = 0; q < 3; ++q) { zmin = std::min(zmin, t.v[3 * q + 2]); zmax = std::max(zmax, t.v[3 * q + 2]); }
    std::vector<double> edges = opt.z_edges;
    if (edges.empty()) {
        const int nb = zmax - zmin > 1e-9 ? std::max(1, opt.max_bands) : 1;
        for (int i = 0; i <= nb; ++i) edges.push_back(zmin + (zmax - zmin) * i / nb);
        if (nb == 1 && !(edges[1] > edges[0])) edges[1] = edges[0] + 1e-6;   // a flat model still needs a prism of some height
    }
    for (size_t i = 1; i < edges.size(); ++i)
        if (!(edges[i] > edges[i - 1])) fail(what, "band boundaries must increase");
    if (edges.size() < 2) fail(what, "at least two band boundaries are needed");
    // Band boundaries usually sit on faces of the model (the top of the base = the bottom of the body), and unit scaling
    // leaves those faces a rounding error beside the boundary: vertices within 1e-9 (relative) of a boundary are put on it
    for (Tri& t : d.out)
        for (int q = 0; q < 3; ++q)
            for (double e : edges)
                if (std::fabs(t.v[3 * q + 2] - e) <= 1e-9 * std::max(1.0, std::fabs(e))) t.v[3 * q + 2] = e;
    std::vector<double> rows;
    for (size_t bnd = 0; bnd + 1 < edges.size(); ++bnd) {
        const double lo = edges[bnd], hi = edges[bnd + 1];
        std::vector<P2> pts;
        for (const Tri& t : d.out) {
            double c[7][3];
            const int n = clip_band(t, lo, hi, c);
            if (n == 0) continue;
            // (vertices were snapped onto the boundaries above)  A piece of a sloping or vertical triangle thinner than eps
            // only grazes the band and is dropped; a horizontal face in the plane shared with the band below is counted there
            // only (it is the top of a solid below, or the bottom of a solid of this band, which adds nothing its walls do not).
            const double eps = 1e-9 * std::max(1.0, std::max(std::fabs(lo), std::fabs(hi)));
            double cz0 = 1e300, cz1 = -1e300, tz0 = 1e300, tz1 = -1e300;
            for (int i = 0; i < n; ++i) { cz0 = std::min(cz0, c[i][2]); cz1 = std::max(cz1, c[i][2]); }
            for (int q = 0; q < 3; ++q) { tz0 = std::min(tz0, t.v[3 * q + 2]); tz1 = std::max(tz1, t.v[3 * q + 2]); }
            const bool flat_triangle = tz1 - tz0 <= eps;
            if (!flat_triangle && cz1 - cz0 <= eps) continue;
            if (flat_triangle && bnd > 0 && tz0 <= lo + eps) continue;
            for (int i = 0; i < n; ++i) pts.push_back(P2{c[i][0], c[i][1]});
        }
        if (pts.empty()) continue;
        std::vector<P2> h = convex_hull(pts);
        if (h.size() < 3) continue;   // a line or a point has no footprint
        if (opt.as_boxes) {
            // minimum-area bounding rectangle: one of its sides lies along an edge of the convex hull (rotating calipers)
            double best = 1e300, r[8] = {0, 0, 1, 0, 0, 0, lo, hi};
            for (size_t i = 0; i < h.size(); ++i) {
                const P2 &a = h[i], &b = h[(i + 1) % h.size()];
                const double len = std::hypot(b.x - a.x, b.y - a.y);
                if (!(len > 0)) continue;
                double ux = (b.x - a.x) / len, uy = (b.y - a.y) / len;
                double u0 = 1e300, u1 = -1e300, v0 = 1e300, v1 = -1e300;
                for (const P2& q : h) {
                    const double pu = q.x * ux + q.y * uy, pv = -q.x * uy + q.y * ux;
                    u0 = std::min(u0, pu); u1 = std::max(u1, pu); v0 = std::min(v0, pv); v1 = std::max(v1, pv);
                }
                const double area = (u1 - u0) * (v1 - v0);
                if (area < best * (1 - 1e-12)) {
                    best = area;
                    const double cu = 0.5 * (u0 + u1), cv = 0.5 * (v0 + v1);
                    r[0] = cu * ux - cv * uy; r[1] = cu * uy + cv * ux;
                    r[2] = ux; r[3] = uy; r[4] = 0.5 * (u1 - u0); r[5] = 0.5 * (v1 - v0);
                }
            }
            // axis in the right half-plane, nearest to +x among the four equivalent ones (a rectangle is its own 90-degree turn)
            for (int turn = 0; turn < 4 && !(r[2] > 0 && std::fabs(r[3]) <= r[2] * (1 + 1e-12)); ++turn) {
                const double c = r[2]; r[2] = -r[3]; r[3] = c; std::swap(r[4], r[5]);
            }
            if (std::fabs(r[3]) < 1e-15) r[3] = 0;
            rows.insert(rows.end(), r, r + 8);
            continue;
        }
        while (int(h.size()) > opt.max_vertices)
            if (!drop_cheapest_edge(h)) fail(what, "cannot simplify a footprint to " + std::to_string(opt.max_vertices) + " vertices");
        double r[19] = {double(h.size()), lo, hi};
        for (size_t i = 0; i < h.size(); ++i) { r[3 + 2 * i] = h[i].x; r[4 + 2 * i] = h[i].y; }
        rows.insert(rows.end(), r, r + 19);
    }
    if (rows.empty()) fail(what, "no band of the model has a footprint");
    return rows;
}

std::vector<double> load_collada_hulls(const std::string& path, const ColladaOptions& opt) {
    const std::string text = ends_with_ci(path, ".zae") ? read_zae_document(path) : slurp(path);
    return collada_hulls_from_text(text, opt, path);
}

}  // namespace mcsim
