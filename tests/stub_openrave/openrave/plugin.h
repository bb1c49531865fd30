// TEST INFRASTRUCTURE: stand-in for <openrave/plugin.h> (OpenRAVE 0.9 is not in this image) declaring only the API
// csrc/openrave_adapter.cpp uses, as remembered from OpenRAVE's public headers, so that the adapter is at least
// compile-checked.  It proves nothing about linking against a real OpenRAVE.
#pragma once
#include <cstdlib>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>
#include <boost/bind.hpp>
#define OPENRAVE_PLUGIN_API extern "C"
namespace OpenRAVE {
typedef double dReal;
struct Vector { dReal x = 0, y = 0, z = 0, w = 0; };
struct Transform {
    Vector rot, trans;  // rot: quaternion with x = cos(a/2), w = sin(a/2) z-component (OpenRAVE stores (cos, sin*axis))
    Transform operator*(const Transform& o) const { (void)o; return *this; }
    Vector operator*(const Vector& v) const { return v; }
    Transform inverse() const { return *this; }
};
struct AABB { Vector pos, extents; };
enum GeometryType { GT_None = 0, GT_Box = 1, GT_Sphere = 2, GT_Cylinder = 3, GT_TriMesh = 4 };
enum InterfaceType { PT_Planner = 1, PT_Robot = 2, PT_Module = 8 };
struct EnvironmentMutex : std::recursive_mutex { typedef std::unique_lock<std::recursive_mutex> scoped_lock; };
class KinBody {
  public:
    class Link {
      public:
        class Geometry {
          public:
            GeometryType GetType() const { return GT_Box; }
            Transform GetTransform() const { return Transform(); }
            Vector GetBoxExtents() const { return Vector(); }
        };
        typedef std::shared_ptr<Geometry> GeometryPtr;
        const std::vector<GeometryPtr>& GetGeometries() const { return geoms_; }
        Transform GetTransform() const { return Transform(); }
        AABB ComputeLocalAABB() const { return AABB(); }
        std::vector<GeometryPtr> geoms_;
    };
    typedef std::shared_ptr<Link> LinkPtr;
    virtual ~KinBody() {}
    virtual bool IsRobot() const { return false; }
    const std::vector<LinkPtr>& GetLinks() const { return links_; }
    Transform GetTransform() const { return Transform(); }
    std::vector<LinkPtr> links_;
};
typedef std::shared_ptr<KinBody> KinBodyPtr;
class RobotBase : public KinBody { public: bool IsRobot() const override { return true; } };
typedef std::shared_ptr<RobotBase> RobotBasePtr;
class EnvironmentBase {
  public:
    EnvironmentMutex& GetMutex() { return m_; }
    void GetRobots(std::vector<RobotBasePtr>& r) { r.clear(); }
    void GetBodies(std::vector<KinBodyPtr>& b) { b.clear(); }
    EnvironmentMutex m_;
};
typedef std::shared_ptr<EnvironmentBase> EnvironmentBasePtr;
class InterfaceBase {
  public:
    typedef std::function<bool(std::ostream&, std::istream&)> InterfaceCommandFn;
    virtual ~InterfaceBase() {}
    void RegisterCommand(const std::string& name, InterfaceCommandFn fn, const std::string& help) { cmds_[name] = fn; (void)help; }
    std::map<std::string, InterfaceCommandFn> cmds_;
};
typedef std::shared_ptr<InterfaceBase> InterfaceBasePtr;
class ModuleBase : public InterfaceBase { public: explicit ModuleBase(EnvironmentBasePtr) {} };
struct PLUGININFO { std::map<InterfaceType, std::vector<std::string> > interfacenames; };
}  // namespace OpenRAVE
