// rt_internal.hpp — shared between the host-only and the CUDA halves of the library.
#pragma once
#include <string>
#include "../../include/rt_b200.h"
#include "host_scene.hpp"

void rt_set_error(const std::string& message);            // thread-local, read by rt_last_error()
namespace rtb { rt_compiled* compile_scene(const Scene& scene, const Camera& cam); }
