// renderer.hpp — C++ host-side mirror of the reference's `Renderer` for the compute path only,
// written above the C ABI (include/ptb200.h).  The reference host is compiled code (C#,
// RayTracing/Graphics/Renderer.cs); with no .NET toolchain here this header is the compiled-host
// stand-in: same member names, argument meaning and error behaviour (exceptions with the native
// message, like `throw new Exception(...)` at Renderer.cs:115-118, 1036-1039).
//
//   Renderer(Window)          Renderer.cs:59-64
//   Init()                    Renderer.cs:66-84     (compute-relevant subset, same order)
//   CreateResources()         Renderer.cs:105-196
//   CreateComputePipeline()   Renderer.cs:293-403
//   ComputeFrame(delta)       Renderer.cs:1006-1040
//   WaitForFences()           Renderer.cs:972
//   Render(delta)             Renderer.cs:933-1004  (headless present-equivalent instead of DrawFrame)
//   Dispose()                 Renderer.cs:1192-1216
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/ptb200.h"

namespace ptb_host {

struct Window {                       // stand-in for Silk.NET IWindow: only FramebufferSize is read
    uint32_t width = 1920, height = 1080;   // App.cs:27
};

class Renderer {
public:
    // seed: RNG seed of jittered samples (no reference counterpart: Test.hlsl is deterministic)
    explicit Renderer(Window w = Window(), int device = 0, uint32_t seed = 0) : window_(w), seed_(seed), device_(device) {}
    ~Renderer() { Dispose(); }
    Renderer(const Renderer &) = delete;
    Renderer &operator=(const Renderer &) = delete;

    void Init()
    {
        CreateResources();
        CreateComputePipeline();
    }

    void CreateResources()
    {
        ptb_config cfg;
        check(nullptr, ptb_default_config(&cfg, window_.width, window_.height));
        cfg.device = device_;
        cfg.seed = seed_;
        check(nullptr, ptb_create(&cfg, &ctx_));
    }

    void CreateComputePipeline()
    {
        if (!ctx_) throw std::runtime_error("CreateComputePipeline before CreateResources");
        pipeline_ready_ = true;       // the kernel is linked into libptb200.so; nothing to load
    }

    void Update(float) {}             // Renderer.cs:86-89: empty in the reference

    void ComputeFrame(float /*delta: unused in the reference too*/)
    {
        if (!pipeline_ready_) throw std::runtime_error("ComputeFrame before CreateComputePipeline");
        check(ctx_, ptb_compute_frame(ctx_));
    }

    void WaitForFences() { check(ctx_, ptb_wait(ctx_)); }

    void Render(float delta)
    {
        ComputeFrame(delta);
        WaitForFences();
        check(ctx_, ptb_present(ctx_, window_.width, window_.height, PTB_RESOLVE_SRGB));
    }

    std::vector<uint8_t> ReadImage()        // RGBA8, tightly packed
    {
        std::vector<uint8_t> img((size_t)window_.width * window_.height * 4);
        check(ctx_, ptb_read_rgba8(ctx_, img.data(), 0));
        return img;
    }

    std::vector<uint8_t> ReadPresented()    // BGRA8 (sRGB-encoded), what the swapchain image would hold
    {
        std::vector<uint8_t> img((size_t)window_.width * window_.height * 4);
        check(ctx_, ptb_read_present(ctx_, img.data(), 0));
        return img;
    }

    ptb_context *Handle() const { return ctx_; }

    void Dispose()
    {
        if (ctx_) {
            ptb_destroy(ctx_);
            ctx_ = nullptr;
        }
        pipeline_ready_ = false;
    }

private:
    static void check(ptb_context *ctx, int status)
    {
        if (status != PTB_OK)
            throw std::runtime_error("ptb status " + std::to_string(status) + ": " + ptb_last_error(ctx));
    }

    Window window_;
    uint32_t seed_ = 0;
    int device_;
    ptb_context *ctx_ = nullptr;
    bool pipeline_ready_ = false;
};

}  // namespace ptb_host
