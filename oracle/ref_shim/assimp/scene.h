// Empty stand-in (see Importer.hpp). Test infrastructure (oracle/_ref build) only.
#pragma once
struct aiMesh; struct aiScene;
