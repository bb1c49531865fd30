// mcsim_capi.cpp -- include/mcsim.h over mcsim::MCModule.
#include "mcsim.h"

#include <cstring>
#include <memory>
#include <sstream>
#include <string>

#include "mc_module.h"

struct mcsim_handle {
    std::unique_ptr<mcsim::MCModule> mod;
    std::string err;
    std::string last;   // reply of the last successful command (mcsim_last_output)
};

static thread_local std::string g_create_err;

static mcsim_handle* make(const mcsim::Scene& sc, int device) {
    auto h = std::make_unique<mcsim_handle>();
    h->mod = std::make_unique<mcsim::MCModule>(sc, device);
    return h.release();
}

extern "C" mcsim_handle* mcsim_create(const char* env_xml_path, const char* fixture_dirs, int device) {
    try {
        if (!env_xml_path) throw std::runtime_error("env_xml_path is NULL");
        std::vector<std::string> dirs;
        if (fixture_dirs) {
            std::istringstream is(fixture_dirs);
            std::string d;
            while (std::getline(is, d, ':')) if (!d.empty()) dirs.push_back(d);
        }
        return make(mcsim::load_scene_xml(env_xml_path, dirs), device);
    } catch (const std::exception& e) {
        g_create_err = e.what();
        return nullptr;
    }
}

extern "C" mcsim_handle* mcsim_create_from_arrays(const double* boxes, int n_boxes, const double* hulls, int n_hulls, int device) {
    try {
        if (n_boxes < 0 || n_hulls < 0 || (n_boxes && !boxes) || (n_hulls && !hulls)) throw std::runtime_error("bad geometry arrays");
        mcsim::Scene sc;
        sc.boxes.assign(boxes, boxes + size_t(n_boxes) * 8);
        sc.hulls.assign(hulls, hulls + size_t(n_hulls) * 8);
        return make(sc, device);
    } catch (const std::exception& e) {
        g_create_err = e.what();
        return nullptr;
    }
}

extern "C" int mcsim_send_command(mcsim_handle* h, const char* cmd, char* out, size_t out_cap, size_t* out_len) {
    if (out && out_cap) out[0] = '\0';
    if (out_len) *out_len = 0;
    if (!h || !cmd) return 1;
    std::string text;
    const bool ok = h->mod->SendCommand(std::string(cmd), text);
    if (!ok) { h->err = h->mod->last_error(); return 1; }
    h->err.clear();
    h->last = text;
    if (out_len) *out_len = text.size();
    if (out && out_cap) {
        const size_t n = text.size() < out_cap - 1 ? text.size() : out_cap - 1;
        std::memcpy(out, text.data(), n);
        out[n] = '\0';
    }
    return 0;
}

extern "C" size_t mcsim_last_output(mcsim_handle* h, char* out, size_t out_cap) {
    if (!h) return 0;
    if (out && out_cap) {
        const size_t n = h->last.size() < out_cap - 1 ? h->last.size() : out_cap - 1;
        std::memcpy(out, h->last.data(), n);
        out[n] = '\0';
    }
    return h->last.size();
}

extern "C" const char* mcsim_last_error(mcsim_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

static int copy_rows(const std::vector<double>& rows, double* out, int cap_rows) {
    const int n = int(rows.size() / 8);
    if (out && cap_rows > 0) std::memcpy(out, rows.data(), sizeof(double) * 8 * size_t(n < cap_rows ? n : cap_rows));
    return n;
}
extern "C" int mcsim_get_boxes(mcsim_handle* h, double* out, int cap_rows) { return h ? copy_rows(h->mod->scene().boxes, out, cap_rows) : 0; }
extern "C" int mcsim_get_hulls(mcsim_handle* h, double* out, int cap_rows) { return h ? copy_rows(h->mod->scene().hulls, out, cap_rows) : 0; }
extern "C" int mcsim_get_poly_hulls(mcsim_handle* h, double* out, int cap_rows) {
    if (!h) return 0;
    const std::vector<double>& rows = h->mod->scene().poly_hulls;
    const int n = int(rows.size() / 19);
    if (out && cap_rows > 0) std::memcpy(out, rows.data(), sizeof(double) * 19 * size_t(n < cap_rows ? n : cap_rows));
    return n;
}
extern "C" int mcsim_get_robot_pose(mcsim_handle* h, double pose[3]) {
    if (!h || !h->mod->scene().has_robot) return 0;
    std::memcpy(pose, h->mod->scene().robot_pose, sizeof(double) * 3);
    return 1;
}
extern "C" void mcsim_io_bytes(mcsim_handle* h, uint64_t* h2d, uint64_t* d2h) {
    if (h) h->mod->io_bytes(h2d, d2h);
}
extern "C" void mcsim_destroy(mcsim_handle* h) { delete h; }
