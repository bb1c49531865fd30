// TEST INFRASTRUCTURE: stand-in for <boost/bind.hpp> (Boost is not in this image) -- just enough for the
// compile check of csrc/openrave_adapter.cpp.
#pragma once
#include <functional>
#include <memory>
namespace boost {
template <class T> using shared_ptr = std::shared_ptr<T>;
using std::bind;
}
using namespace std::placeholders;
