// TEST INFRASTRUCTURE ONLY (oracle/): stand-in for <openrave/plugin.h>.
//
// OpenRAVE is not installed in this image.  This header declares exactly the
// OpenRAVE symbols the *unmodified* reference estimator touches
// (mcsimplugin/MCSimulator.h:132-136,139,150-155,272-282) so that the reference headers compile
// from where they lie under /root/reference.  EnvironmentBase::CheckCollision is served by the
// CPU SAT checker of oracle/sat_cpu.h over the same box/hull table the CUDA path uses
// (the reference's ODE checker + PR2 mesh are not in the repo).
#ifndef ORACLE_STUB_OPENRAVE_PLUGIN_H
#define ORACLE_STUB_OPENRAVE_PLUGIN_H

#include <memory>
#include <vector>
#include <string>
#include <iostream>

namespace OpenRAVE {

typedef double dReal;

struct EnvironmentMutex {
    void lock() {}
    void unlock() {}
};

class RobotBase {
public:
    double pose[3];
    RobotBase() { pose[0] = pose[1] = pose[2] = 0.0; }
    void SetActiveDOFValues(const std::vector<dReal>& v) {
        pose[0] = v[0]; pose[1] = v[1]; pose[2] = v[2];
    }
};
typedef std::shared_ptr<RobotBase> RobotBasePtr;

// Hook implemented by the oracle driver: robot pose -> collides?  (also logs clearance)
typedef bool (*oracle_check_fn)(void* ctx, const double pose[3]);

class EnvironmentBase {
public:
    EnvironmentMutex mtx;
    std::vector<RobotBasePtr> robots;
    oracle_check_fn check;
    void* check_ctx;
    EnvironmentBase() : check(nullptr), check_ctx(nullptr) {
        robots.push_back(RobotBasePtr(new RobotBase()));
    }
    EnvironmentMutex& GetMutex() { return mtx; }
    void GetRobots(std::vector<RobotBasePtr>& out) { out = robots; }
    bool CheckCollision(RobotBasePtr r) { return check(check_ctx, r->pose); }
};
typedef std::shared_ptr<EnvironmentBase> EnvironmentBasePtr;

}  // namespace OpenRAVE

#endif
