// Linux stand-in for the reference's pch.h (which pulls <windows.h>): forced-included when
// compiling /root/reference/RayTracing/BVH.cpp verbatim. Test infrastructure only.
#pragma once
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <iostream>
#include <memory>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>
#define QUALIFIERS static inline
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include "glm/glm.hpp"
#include "glm/gtc/matrix_transform.hpp"
#include "glm/gtc/quaternion.hpp"
