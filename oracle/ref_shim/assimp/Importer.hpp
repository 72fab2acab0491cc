// Empty stand-in so the reference's BVH.cpp (which includes utils/AssimpLoader.h) compiles
// without Assimp. Only the type name is needed. Test infrastructure (oracle/_ref build) only.
#pragma once
namespace Assimp { class Importer {}; }
