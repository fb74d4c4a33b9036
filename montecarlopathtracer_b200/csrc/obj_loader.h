// OBJ/MTL ingest with the reference's two reader semantics:
//   MCPT_LOADER_TINYOBJ   -- tinyobjloader 1.1.1 as QuinModel uses it (Structure.hpp:9-12)
//   MCPT_LOADER_OBJREADER -- PW::FileReader::ObjModel (CVMCTracer/.../ObjReader.cpp:8-259)
#pragma once
#include <string>

#include "host_scene.h"

namespace mcpt {

bool load_obj_scene(const std::string& obj_path, const std::string& mtl_base, int loader, HostScene& out, std::string& err);

}  // namespace mcpt
