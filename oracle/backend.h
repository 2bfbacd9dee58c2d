// oracle/backend.h — TEST INFRASTRUCTURE, never part of the product path.
//
// Selects where the per-ray arithmetic of the CPU oracle comes from:
//   -DNRB_ORACLE_REFERENCE : the reference's own device headers, compiled for the host where
//        they lie under /root/reference/src (mesh_traversal.cuh, mesh_intersection.cuh,
//        disney_brdf.cuh).  Output goes to oracle/_ref/ (git-ignored).  This is the arbiter.
//   otherwise              : oracle/restated/*.h, our plain C++ restatement of the same
//        functions (each citing the reference file:line it follows).  Travels as source.
// Both builds compile the same frame_oracle.cpp (restated kernel bodies) around it.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>

#ifdef NRB_ORACLE_REFERENCE
// vec3.h:5-12 defines __host__/__device__ away when __CUDACC__ is unset; these three lines
// are all the host build needs on top of that (SURVEY.md probe table).
using std::max;
using std::min;
#define __forceinline__ inline
#include "mesh_traversal.cuh"
#include "disney_brdf.cuh"
#include "scene.h"
#include "renderer.h"
#else
#include "restated/geometry.h"
#include "restated/disney.h"
#endif
