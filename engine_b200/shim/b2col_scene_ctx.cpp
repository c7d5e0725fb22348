#include "b2col_scene_ctx.h"

#include <map>
#include <stdexcept>
#include <string>

namespace engine {
namespace {
std::map<Scene*, B2colSceneCtx>& registry()
{
    static std::map<Scene*, B2colSceneCtx> r;  // the engine is single-threaded (reference src/application.cpp:390)
    return r;
}
}  // namespace

B2colSceneCtx* B2colAcquire(Scene* scene)
{
    B2colSceneCtx& e = registry()[scene];
    if (!e.ctx) {
        b2col_config cfg{};
        cfg.use_cuda_graph = 1;  // frames of an unchanged scene shape replay one captured graph
        const b2col_status st = b2col_create(&e.ctx, &cfg);
        if (st != B2COL_OK) {
            registry().erase(scene);
            throw std::runtime_error(std::string("b2col_create: ") + b2col_last_error(nullptr));
        }
    }
    ++e.users;
    return &e;
}

void B2colRelease(Scene* scene)
{
    auto it = registry().find(scene);
    if (it == registry().end()) return;
    if (--it->second.users <= 0) {
        b2col_destroy(it->second.ctx);
        registry().erase(it);
    }
}
}  // namespace engine
