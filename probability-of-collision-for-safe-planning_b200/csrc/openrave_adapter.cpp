// openrave_adapter.cpp -- the OpenRAVE plugin shape of the module (built only where OpenRAVE 0.9 + Boost exist:
// `make openrave`).  Same three entry points and the same interface name as the reference plugin
// (mcsimplugin/mcsimplugin.cpp:236-255: CreateInterfaceValidated / GetPluginAttributesValidated / DestroyPlugin,
// PT_Module "mcmodule"), so `RaveLoadPlugin` + `RaveCreateModule(env, 'MCModule')` of MCSimulation.py:133-135 work
// unchanged.  Every command is forwarded to mcsim::MCModule (mc_module.h), which owns the CUDA context.
//
// The collision world: the reference asks OpenRAVE per pose (MCSimulator.h:269-285).  Here the geometry is read
// ONCE from the loaded environment through the OpenRAVE API -- every box geometry of every non-robot body becomes an
// obstacle prism, every robot link becomes one hull prism (the link's local AABB, in the robot frame) -- unless the
// module is created with an argument string "scene=<env.xml> [fixtures=<dir>]", in which case the XML extractor of
// scene_xml.cpp is used instead.
#include <openrave/plugin.h>
#include <boost/bind.hpp>

#include <cmath>
#include <sstream>

#include "mc_module.h"

using namespace OpenRAVE;

namespace {

double yaw_of(const Transform& t) {  // rotation about +z of a quaternion (w, x, y, z)
    return 2.0 * std::atan2(double(t.rot.w), double(t.rot.x));
}

void push_prism(std::vector<double>& rows, const Transform& t, const Vector& half) {
    const double yaw = yaw_of(t);
    const double r[8] = {double(t.trans.x), double(t.trans.y), std::cos(yaw), std::sin(yaw), double(half.x), double(half.y),
                         double(t.trans.z - half.z), double(t.trans.z + half.z)};
    rows.insert(rows.end(), r, r + 8);
}

mcsim::Scene scene_from_environment(EnvironmentBasePtr penv) {
    mcsim::Scene sc;
    EnvironmentMutex::scoped_lock lock(penv->GetMutex());
    std::vector<RobotBasePtr> robots;
    penv->GetRobots(robots);
    std::vector<KinBodyPtr> bodies;
    penv->GetBodies(bodies);
    for (size_t b = 0; b < bodies.size(); ++b) {
        if (bodies[b]->IsRobot()) continue;
        const std::vector<KinBody::LinkPtr>& links = bodies[b]->GetLinks();
        for (size_t l = 0; l < links.size(); ++l) {
            const std::vector<KinBody::Link::GeometryPtr>& geoms = links[l]->GetGeometries();
            for (size_t g = 0; g < geoms.size(); ++g)
                if (geoms[g]->GetType() == GT_Box)
                    push_prism(sc.boxes, links[l]->GetTransform() * geoms[g]->GetTransform(), geoms[g]->GetBoxExtents());
        }
    }
    if (!robots.empty()) {
        const Transform world_from_robot = robots[0]->GetTransform();
        const Transform robot_from_world = world_from_robot.inverse();
        sc.robot_pose[0] = world_from_robot.trans.x;
        sc.robot_pose[1] = world_from_robot.trans.y;
        sc.robot_pose[2] = yaw_of(world_from_robot);
        sc.has_robot = true;
        const std::vector<KinBody::LinkPtr>& links = robots[0]->GetLinks();
        for (size_t l = 0; l < links.size(); ++l) {
            const AABB box = links[l]->ComputeLocalAABB();  // in the link frame
            Transform t = robot_from_world * links[l]->GetTransform();
            t.trans = t * box.pos;
            t.trans.z += world_from_robot.trans.z;             // hull z is absolute
            push_prism(sc.hulls, t, box.extents);
        }
    }
    return sc;
}

class MCModuleAdapter : public ModuleBase {
  public:
    MCModuleAdapter(EnvironmentBasePtr penv, std::istream& sinput) : ModuleBase(penv) {
        std::string tok, xml;
        std::vector<std::string> fixtures;
        int device = 0;
        while (sinput >> tok) {
            if (tok.compare(0, 6, "scene=") == 0) xml = tok.substr(6);
            else if (tok.compare(0, 9, "fixtures=") == 0) fixtures.push_back(tok.substr(9));
            else if (tok.compare(0, 7, "device=") == 0) device = std::atoi(tok.c_str() + 7);
        }
        impl_.reset(new mcsim::MCModule(xml.empty() ? scene_from_environment(penv) : mcsim::load_scene_xml(xml, fixtures), device));
        // every command of mcsim::MCModule -- the reference's 15 RegisterCommand names (mcsimplugin.cpp:13-44) and this
        // build's additions -- straight from its table, so the two front ends cannot drift apart ("help" is OpenRAVE's own)
        for (const auto& cmd : mcsim::MCModule::command_list()) {
            if (cmd.first == "help") continue;
            RegisterCommand(cmd.first, boost::bind(&MCModuleAdapter::Forward, this, cmd.first, _1, _2), cmd.second);
        }
    }
    virtual ~MCModuleAdapter() {}

    bool Forward(const std::string& name, std::ostream& sout, std::istream& sinput) {
        std::stringstream line;
        line << name << " " << sinput.rdbuf();
        return impl_->SendCommand(sout, line);  // std::runtime_error propagates as an OpenRAVE/Python exception
    }

  private:
    boost::shared_ptr<mcsim::MCModule> impl_;
};

}  // namespace

InterfaceBasePtr CreateInterfaceValidated(InterfaceType type, const std::string& interfacename, std::istream& sinput, EnvironmentBasePtr penv) {
    if (type == PT_Module && interfacename == "mcmodule") return InterfaceBasePtr(new MCModuleAdapter(penv, sinput));
    return InterfaceBasePtr();
}

void GetPluginAttributesValidated(PLUGININFO& info) { info.interfacenames[PT_Module].push_back("MCModule"); }

OPENRAVE_PLUGIN_API void DestroyPlugin() {}
