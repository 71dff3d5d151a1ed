// Link-time stand-ins for Madrona's Vulkan batch renderer (external/madrona/src/render/
// render_mgr.cpp, api.cpp). There is no graphics stack on this box; src/mgr.cpp:41-66
// never constructs these objects when enableBatchRenderer == false, which is the only
// mode gd_ref_create uses. Every entry point that would render aborts loudly.
// Test infrastructure (oracle/_ref) only.
#include <madrona/render/api.hpp>
#include <madrona/render/render_mgr.hpp>
#include <cstdio>
#include <cstdlib>

namespace madrona::render {

struct RenderContext {};

[[noreturn]] static void unavailable(const char *what)
{
    fprintf(stderr, "gd_ref: %s: renderer unavailable on this box\n", what);
    abort();
}

GPUHandle::GPUHandle(APIBackendSelect backend, GPUDevice *dev) : backend_(backend), dev_(dev) {}
GPUHandle::~GPUHandle() {}
APILibHandle::APILibHandle(APIBackendSelect backend, APILib *lib) : backend_(backend), lib_(lib) {}
APILibHandle::~APILibHandle() {}

APILibHandle APIManager::loadDefaultLib(APIBackendSelect) { unavailable("APIManager::loadDefaultLib"); }
APIManager::APIManager(APILib *, const Config &, APIBackendSelect) { unavailable("APIManager"); }
APIManager::APIManager(APIManager &&o) : chosen_backend_(o.chosen_backend_), backend_(o.backend_) { o.backend_ = nullptr; }
APIManager::~APIManager() {}
GPUHandle APIManager::initGPU(CountT) { unavailable("APIManager::initGPU"); }

RenderManager::RenderManager(APIBackend *, GPUDevice *, const Config &) { unavailable("RenderManager"); }
RenderManager::RenderManager(RenderManager &&o) : rctx_(std::move(o.rctx_)) {}
RenderManager::~RenderManager() {}
CountT RenderManager::loadObjects(Span<const imp::SourceObject>, Span<const imp::SourceMaterial>,
                                  Span<const imp::SourceTexture>) { unavailable("loadObjects"); }
void RenderManager::configureLighting(Span<const LightConfig>) { unavailable("configureLighting"); }
const RenderECSBridge *RenderManager::bridge() const { unavailable("bridge"); }
void RenderManager::readECS() { unavailable("readECS"); }
void RenderManager::batchRender() { unavailable("batchRender"); }
const uint8_t *RenderManager::batchRendererRGBOut() const { unavailable("rgb_tensor"); }
const float *RenderManager::batchRendererDepthOut() const { unavailable("depth_tensor"); }

}
