// TEST INFRASTRUCTURE (oracle/_ref build only) -- not product code.
//
// Stand-in for the reference's Win32/D3D precompiled header
// (/root/reference/MCRT/QuinEngine/stdafx.h:1-45) so that the reference's own
// host code -- Utils/KDTree.hpp, Utils/Structure.hpp and the vendored
// 3rdparty/include/tiny_obj_loader.h -- compiles UNMODIFIED with g++ from where
// it lies under /root/reference.  Only the scalar typedefs and two POD stubs the
// headers name are provided; nothing of D3D is emulated.
#pragma once
#include <assert.h>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

typedef float FLOAT;
typedef unsigned int UINT;
typedef int INT;
typedef int BOOL;

// Structure.hpp:186-223 declares GPU-mirrored PODs with these two D3DX types.
// They are never touched by the tree builder; plain aggregates suffice.
struct D3DXVECTOR3 { float x, y, z; };
struct D3DXMATRIX { float m[4][4]; };

#include <tiny_obj_loader.h>
