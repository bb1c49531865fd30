/* mcsim.h -- headless C ABI of the plugin host (libmcsim.so): the mcsimplugin command interface without
 * OpenRAVE.  One handle == one `MCModule` of the reference (mcsimplugin/mcsimplugin.cpp:7-46, created by
 * RaveCreateModule(env, 'MCModule'), MCSimulation.py:133-135) together with the collision world the reference
 * gets from `env.Load(...)` (MCSimulation.py:123-124).
 *
 *   reference call (Python 2, openravepy)                        this ABI
 *   ------------------------------------------------------      ------------------------------------------
 *   env.Load('data/pr2test2.env.xml'); RaveCreateModule(...)     mcsim_create(env_xml, fixture_dirs, device)
 *   MCModule.SendCommand('setAlphas 6.25e-08 ...') -> str|None   mcsim_send_command(h, "setAlphas ...", out, cap)
 *   (module destroyed with the environment)                      mcsim_destroy(h)
 *
 * Commands, token grammar and returned text are those of mcsimplugin.cpp:49-231 (see INTEGRATION.md for the
 * table and for the additional commands).  Plain pointers and sizes; no exceptions cross this boundary.
 * There is no CPU path: mcsim_create fails when no sm_100 device is visible.
 */
#ifndef MCSIM_H
#define MCSIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcsim_handle mcsim_handle;

/* env_xml_path: an OpenRAVE environment XML (obstacle boxes + <Robot>); fixture_dirs: ':'-separated directories
 * searched for files the env names but the repository does not ship (ikeatable.kinbody.xml, the PR2 model as
 * pr2-beta-static.fixture.xml); may be NULL.  device: CUDA ordinal, or -1 for a parse-only handle (setters,
 * queries and scene extraction work; run* commands fail: there is no CPU path).
 * Returns NULL on failure: see mcsim_last_error(NULL). */
mcsim_handle* mcsim_create(const char* env_xml_path, const char* fixture_dirs, int device);

/* The same with the geometry given directly: boxes n_boxes x 8 (world frame), hulls n_hulls x 8 (robot frame),
 * rows = cx, cy, cos(yaw), sin(yaw), hx, hy, zlo, zhi.  n_hulls may be 0 (geometry sent later by setScene /
 * setGeometry). */
mcsim_handle* mcsim_create_from_arrays(const double* boxes, int n_boxes, const double* hulls, int n_hulls, int device);

/* ModuleBase::SendCommand: 0 = the handler returned true and `out` holds its output text (NUL-terminated,
 * truncated to out_cap - 1 characters; the untruncated length is stored in *out_len when out_len != NULL);
 * 1 = unknown command or handler failure (Python's SendCommand would return None or raise): message in
 * mcsim_last_error(h). */
int mcsim_send_command(mcsim_handle* h, const char* cmd, char* out, size_t out_cap, size_t* out_len);

/* The reply of the last successful command again (NUL-terminated, truncated to out_cap - 1 characters); returns its
 * full length.  For replies longer than the buffer given to mcsim_send_command -- fetching them must not re-run the
 * command (runGMMEstimationBatch 10000 returns ~90 KB and advances the run counter). */
size_t mcsim_last_output(mcsim_handle* h, char* out, size_t out_cap);

const char* mcsim_last_error(mcsim_handle* h);

/* The extracted geometry (so a caller can check or reuse it).  Copies up to cap rows of 8 doubles; returns the
 * number of rows available. */
int mcsim_get_boxes(mcsim_handle* h, double* out, int cap_rows);
int mcsim_get_hulls(mcsim_handle* h, double* out, int cap_rows);
/* convex-polygon robot hulls (<Geom type="prism"> of the robot file, or setRobotPolygons): rows of 19 doubles =
 * n, zlo, zhi, x0, y0, ..., x7, y7 (counter-clockwise, robot frame) */
int mcsim_get_poly_hulls(mcsim_handle* h, double* out, int cap_rows);
/* x, y, yaw of the <Robot> element; returns 0 if the scene had none */
int mcsim_get_robot_pose(mcsim_handle* h, double pose[3]);

/* bytes copied host->device / device->host by the last run* command */
void mcsim_io_bytes(mcsim_handle* h, uint64_t* h2d, uint64_t* d2h);

void mcsim_destroy(mcsim_handle* h);

#ifdef __cplusplus
}
#endif
#endif /* MCSIM_H */
