// TEST INFRASTRUCTURE ONLY — stand-in for <zmqpp.hpp> so that the reference's NetworkPlot.h (plotting/NetworkPlot.h:9,
// 53-54) compiles on a box without libzmq: the reference wrappers only hold pointers to these types.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>
namespace zmqpp {
class socket;
class context;
class message;
}  // namespace zmqpp
