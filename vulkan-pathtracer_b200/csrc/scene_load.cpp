// placeholder — replaced by the glTF loader
#include <string>
#include "../../include/b200pt.h"
namespace pt { void set_error(const std::string& msg); }
extern "C" {
int pt_scene_load(const char*, PtScene**) { pt::set_error("pt_scene_load: not built yet"); return PT_ERR_IO; }
int pt_scene_desc(const PtScene*, PtSceneDesc*) { return PT_ERR_INVALID_ARGUMENT; }
void pt_scene_free(PtScene*) {}
}
