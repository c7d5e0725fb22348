#include "b2col_scene_ctx.h"

#include <cstdlib>
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

void B2colUseDevices(B2colSceneCtx* shared, const int* devices, int n)
{
    if (!shared->ranks.empty() || n < 2) return;
    std::vector<b2col_ctx*> ranks{shared->ctx};
    auto undo = [&](const std::string& why) {
        for (std::size_t i = 1; i < ranks.size(); ++i) b2col_destroy(ranks[i]);
        throw std::runtime_error(why);
    };
    for (int i = 1; i < n; ++i) {  // devices[0] is the building context's
        b2col_config cfg{};
        cfg.device = devices[i];
        b2col_ctx* c = nullptr;
        if (b2col_create(&c, &cfg) != B2COL_OK) undo(std::string("b2col_create (replica): ") + b2col_last_error(nullptr));
        ranks.push_back(c);
    }
    if (b2col_comm_init_all(ranks.data(), n) != B2COL_OK) undo(std::string("b2col_comm_init_all: ") + b2col_last_error(shared->ctx));
    shared->ranks = ranks;
    shared->replicas_fresh = false;
}

namespace {
// B2COL_DEVICES -> device list ("" / unset / one device: single GPU)
std::vector<int> DevicesFromEnv()
{
    std::vector<int> out;
    const char* e = std::getenv("B2COL_DEVICES");
    if (!e || !*e) return out;
    const std::string s(e);
    if (s == "all" || s.find(',') == std::string::npos) {
        int have = 0;
        const int want = s == "all" ? 1 << 30 : std::atoi(s.c_str());
        b2col_device_count(&have);
        for (int d = 0; d < have && d < want; ++d) out.push_back(d);
        return out;
    }
    std::size_t at = 0;
    while (at <= s.size()) {
        const std::size_t comma = s.find(',', at);
        const std::string tok = s.substr(at, comma == std::string::npos ? std::string::npos : comma - at);
        if (!tok.empty()) out.push_back(std::atoi(tok.c_str()));
        if (comma == std::string::npos) break;
        at = comma + 1;
    }
    return out;
}
}  // namespace

B2colSceneCtx* B2colAcquire(Scene* scene)
{
    B2colSceneCtx& e = registry()[scene];
    if (!e.ctx) {
        const std::vector<int> devices = DevicesFromEnv();
        b2col_config cfg{};
        cfg.use_cuda_graph = 1;  // frames of an unchanged scene shape replay one captured graph
        if (!devices.empty()) cfg.device = devices[0];
        const b2col_status st = b2col_create(&e.ctx, &cfg);
        if (st != B2COL_OK) {
            registry().erase(scene);
            throw std::runtime_error(std::string("b2col_create: ") + b2col_last_error(nullptr));
        }
        if (devices.size() > 1) {
            try {
                B2colUseDevices(&e, devices.data(), static_cast<int>(devices.size()));
            }
            catch (...) {
                b2col_destroy(e.ctx);
                registry().erase(scene);
                throw;
            }
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
        for (std::size_t i = 1; i < it->second.ranks.size(); ++i) b2col_destroy(it->second.ranks[i]);  // each finalizes its communicator
        b2col_destroy(it->second.ctx);
        registry().erase(it);
    }
}
}  // namespace engine
