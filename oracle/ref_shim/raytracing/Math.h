// Case shim: utils/DebugDraw.h:3 includes "raytracing/Math.h" (Windows is case-insensitive).
#pragma once
#include "RayTracing/Math.h"
