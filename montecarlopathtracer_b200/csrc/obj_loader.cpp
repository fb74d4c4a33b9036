#include "obj_loader.h"

namespace mcpt {

bool load_obj_scene(const std::string&, const std::string&, int, HostScene&, std::string& err) {
    err = "OBJ loader not built yet";
    return false;
}

}  // namespace mcpt
