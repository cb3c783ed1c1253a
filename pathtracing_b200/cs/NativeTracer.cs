// NativeTracer.cs — the P/Invoke binding a maintainer of chairclr/PathTracing adds next to
// RayTracing/Graphics/Renderer.cs to run the tracing path on libptb200.so instead of the Vulkan
// compute pipeline.  SOURCE ONLY: there is no .NET SDK on the build or GPU box, so this file is
// not compiled or exercised here (the executable stand-in is pathtracing_b200/tracer.py, which
// makes the same calls in the same order).  EVERY entry point of include/ptb200.h has a [DllImport] here
// (tests/test_abi.py diffs the two name lists); the public members wrap the ones a renderer needs.
//
// Replaces, in Renderer.cs: CreateResources (105-196) + CreateComputePipeline (293-403) ->
// NativeTracer ctor; ComputeFrame (1006-1040) -> ComputeFrame(); ResetFences/WaitForFences
// (970-972) -> WaitForFences(); the sampled ComputeImage (1086-1089) -> ReadImage()/Present().
using System;
using System.Runtime.InteropServices;

namespace RayTracing.Graphics;

public sealed unsafe class NativeTracer : IDisposable
{
    private const string Lib = "ptb200";   // libptb200.so / ptb200.dll on the loader path

    [StructLayout(LayoutKind.Sequential)]
    public struct Config
    {
        public uint StructSize;
        public int Device;
        public uint Width, Height;
        public float UvDivX, UvDivY;            // Test.hlsl:6 `resolution`
        public fixed float Cam[3];              // Test.hlsl:8
        public fixed float Centre[3];           // Test.hlsl:12
        public float Radius;                    // Test.hlsl:13
        public uint Seed, MaxDepth, Flags;
        public IntPtr Stream;
    }

    [DllImport(Lib)] private static extern int ptb_default_config(Config* cfg, uint width, uint height);
    [DllImport(Lib)] private static extern int ptb_create(Config* cfg, out IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_destroy(IntPtr ctx);
    [DllImport(Lib)] private static extern IntPtr ptb_last_error(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_compute_frame(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_wait(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_read_rgba8(IntPtr ctx, void* hostDst, nuint pitchBytes);
    [DllImport(Lib)] private static extern int ptb_read_rgba8_async(IntPtr ctx, void* hostDst, nuint pitchBytes);
    [DllImport(Lib)] private static extern int ptb_accum_reset(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_render_samples(IntPtr ctx, uint firstSample, uint nSamples, int jitter);
    [DllImport(Lib)] private static extern int ptb_resolve(IntPtr ctx, uint flags);
    [DllImport(Lib)] private static extern int ptb_render_resolve(IntPtr ctx, uint firstSample, uint nSamples, int jitter, uint flags);
    [DllImport(Lib)] private static extern uint ptb_accum_spp(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_present(IntPtr ctx, uint dstW, uint dstH, uint flags);
    [DllImport(Lib)] private static extern int ptb_read_present(IntPtr ctx, void* hostDst, nuint pitchBytes);

    // ---- scene path (no reference counterpart: the reference's scene is the literals of Test.hlsl:8,12-13)
    [StructLayout(LayoutKind.Sequential)]
    public struct Material                      // 32 bytes == ptb_material
    {
        public uint Type;                       // 0 Lambert, 1 metal, 2 dielectric
        public fixed float Albedo[3];
        public float Fuzz, Ior;
        public fixed float Pad[2];
    }

    [StructLayout(LayoutKind.Sequential)]
    public struct SceneDesc                     // == ptb_scene_desc
    {
        public uint StructSize, NSpheres;
        public float* Spheres;                  // n x (cx,cy,cz,r)
        public uint* SphereMaterial;
        public uint NTriangles, NMaterials;
        public float* TriVertices;              // n x 9
        public uint* TriMaterial;
        public Material* Materials;
        public fixed float CamOrigin[3];        // Test.hlsl:8
        public fixed float CamU[3];             // Test.hlsl:10 generalised: ray = normalize(uv.x*U + uv.y*V + W)
        public fixed float CamV[3];
        public fixed float CamW[3];
        public fixed float SkyHorizon[3];
        public fixed float SkyZenith[3];
        public float TMin;
        public uint ShadeMode, MaxLeaf, Flags;
    }

    [DllImport(Lib)] private static extern int ptb_upload_scene(IntPtr ctx, SceneDesc* desc);
    [DllImport(Lib)] private static extern int ptb_clear_scene(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_scene_set_wave_samples(IntPtr ctx, uint samplesPerWave);
    [DllImport(Lib)] private static extern int ptb_scene_set_sun(IntPtr ctx, float[] dirToLight, float[] radiance);
    [DllImport(Lib)] private static extern int ptb_trace_rays(IntPtr ctx, ulong nRays, float* origins, float* dirs, float tMin,
                                                              float tMax, int mode, int* outPrim, float* outT);
    [DllImport(Lib)] private static extern int ptb_trace_rays_static(IntPtr ctx, ulong nRays, float* origins, float* dirs, float tMin,
                                                                     float tMax, int mode, int* outPrim, float* outT);

    // ---- multi-GPU: row shards, external accumulation buffers, fused merge over NVLink peer memory
    [DllImport(Lib)] private static extern int ptb_set_row_shard_bands(IntPtr ctx, uint rowOffset, uint rowStride, uint rowsPerBand);
    [DllImport(Lib)] private static extern int ptb_set_accum_buffer(IntPtr ctx, void* devicePtr);
    [DllImport(Lib)] private static extern int ptb_set_accum_spp(IntPtr ctx, uint sppTotal);
    [DllImport(Lib)] private static extern int ptb_accum_clear(IntPtr ctx);
    [DllImport(Lib)] private static extern void* ptb_device_accum(IntPtr ctx);
    [DllImport(Lib)] private static extern void* ptb_device_rgba8(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_ipc_export(IntPtr ctx, int which, byte* handle64);
    [DllImport(Lib)] private static extern int ptb_ipc_open(IntPtr ctx, byte* handle64, out IntPtr devicePtr);
    [DllImport(Lib)] private static extern int ptb_peer_merge(IntPtr ctx, uint nPeers, IntPtr* peerAccum, IntPtr dstImage,
                                                              uint slabFirst, uint slabCount, uint sppTotal, uint flags);

    // ---- the rest of include/ptb200.h: versions, planes, checkpoint, multi-frame launch, peer group, inspection, timing
    [StructLayout(LayoutKind.Sequential)]
    public struct SceneInfo                     // == ptb_scene_info
    {
        public uint NSpheres, NTriangles, SphereNodes, TriangleNodes, SphereLeaves, TriangleLeaves, MaxLeaf;
        public float BuildMs;
        public ulong DeviceBytes;
    }

    [StructLayout(LayoutKind.Sequential)]
    public struct SceneStats                    // == ptb_scene_stats
    {
        public ulong Rays, NodesVisited, PrimsTested, Scatters, Paths, MaxStack;
    }

    [DllImport(Lib)] private static extern int ptb_abi_version();
    [DllImport(Lib)] private static extern int ptb_device_count();
    [DllImport(Lib)] private static extern int ptb_kernel_mode(IntPtr ctx);
    [DllImport(Lib)] private static extern ulong ptb_launch_count(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_compute_frames(IntPtr ctx, uint nFrames);
    [DllImport(Lib)] private static extern int ptb_read_f32(IntPtr ctx, float* hostDst, nuint pitchBytes);
    [DllImport(Lib)] private static extern int ptb_read_hit_ids(IntPtr ctx, int* hostDst, nuint pitchBytes);
    [DllImport(Lib)] private static extern void* ptb_device_f32(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_read_accum(IntPtr ctx, float* hostDst, nuint pitchBytes);
    [DllImport(Lib)] private static extern int ptb_write_accum(IntPtr ctx, float* hostSrc, nuint pitchBytes, uint sppTotal);
    [DllImport(Lib)] private static extern int ptb_set_row_shard(IntPtr ctx, uint rowOffset, uint rowStride);
    [DllImport(Lib)] private static extern int ptb_scene_info_get(IntPtr ctx, SceneInfo* info);
    [DllImport(Lib)] private static extern int ptb_scene_stats_get(IntPtr ctx, SceneStats* stats);
    [DllImport(Lib)] private static extern int ptb_scene_stats_reset(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_scene_queue_sizes(IntPtr ctx, uint* sizes, uint cap, out uint n);
    [DllImport(Lib)] private static extern int ptb_scene_read_bvh(IntPtr ctx, int which, void* nodes, uint capNodes, out uint nNodes,
                                                                  uint* order, uint capPrims, out uint nPrims);
    [DllImport(Lib)] private static extern void* ptb_device_peer_flags(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_peer_barrier(IntPtr ctx, uint nPeers, IntPtr* peerFlags, uint selfRank, uint phase, uint epoch);
    [DllImport(Lib)] private static extern int ptb_peer_group_create(IntPtr ctx);
    [DllImport(Lib)] private static extern void* ptb_peer_group_local(IntPtr ctx, int which);
    [DllImport(Lib)] private static extern int ptb_peer_group_bind(IntPtr ctx, uint nPeers, uint selfRank, uint dstRank, IntPtr* partial0,
                                                                   IntPtr* partial1, IntPtr* flags, IntPtr dstImage0, IntPtr dstImage1);
    [DllImport(Lib)] private static extern int ptb_peer_frame(IntPtr ctx, uint firstSample, uint nSamples, int jitter, uint sppTotal, uint flags);
    [DllImport(Lib)] private static extern int ptb_peer_finish(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_peer_read_total(IntPtr ctx, float* hostDst, out uint slabFirst, out uint slabCount);
    [DllImport(Lib)] private static extern int ptb_timer_begin(IntPtr ctx);
    [DllImport(Lib)] private static extern int ptb_timer_end(IntPtr ctx, out float elapsedMs);
    [DllImport(Lib)] private static extern int ptb_selftest_exact_math(IntPtr ctx, int which, float divisor, out ulong mismatches, out ulong tested);

    public const uint ResolveSrgb = 1, ResolveBgra = 2, PeerRenderOnly = 0x20;
    public const uint CfgKeepF32 = 1, CfgKeepHitIds = 2, CfgForceIeee = 8, CfgSpirvArith = 0x10;

    private IntPtr _ctx;
    public uint Width { get; }
    public uint Height { get; }

    // ≙ CreateResources + CreateComputePipeline: size from Window.FramebufferSize (Renderer.cs:128-129).
    public NativeTracer(uint width, uint height, int device = 0, uint seed = 0, uint maxDepth = 0, uint flags = 0)
    {
        Config cfg;
        Check(IntPtr.Zero, ptb_default_config(&cfg, width, height));
        cfg.Device = device;
        cfg.Seed = seed;
        cfg.MaxDepth = maxDepth;
        cfg.Flags = flags;                      // CfgKeepF32 | CfgKeepHitIds | CfgForceIeee | CfgSpirvArith
        Check(IntPtr.Zero, ptb_create(&cfg, out _ctx));
        Width = width;
        Height = height;
    }

    // The reference throws `new Exception(string)` on any non-Success result (Renderer.cs:1036-1039).
    private static void Check(IntPtr ctx, int status)
    {
        if (status != 0)
            throw new Exception($"ptb status {status}: {Marshal.PtrToStringUTF8(ptb_last_error(ctx))}");
    }

    public void ComputeFrame() => Check(_ctx, ptb_compute_frame(_ctx));                 // Renderer.cs:1006-1040
    public void WaitForFences() => Check(_ctx, ptb_wait(_ctx));                         // Renderer.cs:972
    public void ResetAccumulation() => Check(_ctx, ptb_accum_reset(_ctx));
    public void RenderSamples(uint first, uint count, bool jitter = true) =>
        Check(_ctx, ptb_render_samples(_ctx, first, count, jitter ? 1 : 0));
    public void Resolve(uint flags = 0) => Check(_ctx, ptb_resolve(_ctx, flags));
    /// RenderSamples + Resolve as one kernel launch where possible (literal scene, no row shard, no sRGB); same results.
    public void RenderResolve(uint first, uint count, bool jitter = true, uint flags = 0) =>
        Check(_ctx, ptb_render_resolve(_ctx, first, count, jitter ? 1 : 0, flags));
    public uint AccumulatedSamples => ptb_accum_spp(_ctx);

    public void ReadImage(Span<byte> rgba8)                                              // W*H*4 bytes
    {
        if (rgba8.Length < Width * Height * 4) throw new ArgumentException("buffer too small");
        fixed (byte* p = rgba8) Check(_ctx, ptb_read_rgba8(_ctx, p, 0));
    }

    public void ReadImageAsync(byte* pinned) => Check(_ctx, ptb_read_rgba8_async(_ctx, pinned, 0));

    public void Present(uint dstW, uint dstH, Span<byte> bgra8, bool srgb = true)        // ≙ DrawFrame's blit
    {
        Check(_ctx, ptb_present(_ctx, dstW, dstH, srgb ? ResolveSrgb : 0));
        fixed (byte* p = bgra8) Check(_ctx, ptb_read_present(_ctx, p, 0));
    }

    // Scene upload + GPU BVH build; afterwards ComputeFrame / RenderSamples trace the uploaded scene.
    public void UploadScene(ref SceneDesc desc)
    {
        desc.StructSize = (uint)sizeof(SceneDesc);
        fixed (SceneDesc* p = &desc) Check(_ctx, ptb_upload_scene(_ctx, p));
    }

    public void ClearScene() => Check(_ctx, ptb_clear_scene(_ctx));
    /// Samples of every pixel one wave of the scene path carries (0 = as many as fit the path queues). Scheduling only.
    public void SetWaveSamples(uint samplesPerWave) => Check(_ctx, ptb_scene_set_wave_samples(_ctx, samplesPerWave));
    /// <summary>Directional light sampled by next-event estimation (any-hit shadow rays) at Lambert hits; radiance 0 = off.</summary>
    public void SetSun(float[] dirToLight, float[] radiance) => Check(_ctx, ptb_scene_set_sun(_ctx, dirToLight, radiance));

    public void TraceRays(ReadOnlySpan<float> origins, ReadOnlySpan<float> dirs, float tMin, float tMax, bool anyHit,
                          Span<int> prim, Span<float> t)
    {
        fixed (float* o = origins, d = dirs, tt = t)
        fixed (int* pp = prim)
            Check(_ctx, ptb_trace_rays(_ctx, (ulong)prim.Length, o, d, tMin, tMax, anyHit ? 1 : 0, pp, anyHit ? null : tt));
    }

    /// ptb_trace_rays_static: same results with the device staging in the library's static array (no device allocation in the
    /// call); at most 65,536 rays (PTB_TRACE_RAYS_STATIC_MAX), calls on one GPU must not overlap.
    public void TraceRaysStatic(ReadOnlySpan<float> origins, ReadOnlySpan<float> dirs, float tMin, float tMax, bool anyHit,
                                Span<int> prim, Span<float> t)
    {
        fixed (float* o = origins, d = dirs, tt = t)
        fixed (int* pp = prim)
            Check(_ctx, ptb_trace_rays_static(_ctx, (ulong)prim.Length, o, d, tMin, tMax, anyHit ? 1 : 0, pp, anyHit ? null : tt));
    }

    public void SetRowShard(uint rowOffset, uint rowStride, uint rowsPerBand = 1) =>
        Check(_ctx, ptb_set_row_shard_bands(_ctx, rowOffset, rowStride, rowsPerBand));
    public IntPtr DeviceAccum => (IntPtr)ptb_device_accum(_ctx);
    public IntPtr DeviceImage => (IntPtr)ptb_device_rgba8(_ctx);
    public void ClearPartials() => Check(_ctx, ptb_accum_clear(_ctx));

    public byte[] IpcExport(int which)                                   // 0 accumulation buffer, 1 RGBA8 image
    {
        var h = new byte[64];
        fixed (byte* p = h) Check(_ctx, ptb_ipc_export(_ctx, which, p));
        return h;
    }

    public IntPtr IpcOpen(byte[] handle)
    {
        fixed (byte* p = handle) { Check(_ctx, ptb_ipc_open(_ctx, p, out IntPtr ptr)); return ptr; }
    }

    // This GPU's pixel slab: sum the peers' partial buffers in rank order, resolve, store into dstImage (peer memory).
    public void PeerMerge(ReadOnlySpan<IntPtr> peerAccum, IntPtr dstImage, uint slabFirst, uint slabCount, uint sppTotal, uint flags = 0)
    {
        fixed (IntPtr* p = peerAccum)
            Check(_ctx, ptb_peer_merge(_ctx, (uint)peerAccum.Length, p, dstImage, slabFirst, slabCount, sppTotal, flags));
    }

    // ---- members over the remaining entry points
    public static int AbiVersion => ptb_abi_version();
    public static int DeviceCount => ptb_device_count();
    public int KernelMode => ptb_kernel_mode(_ctx);
    public ulong LaunchCount => ptb_launch_count(_ctx);
    /// n reference frames in ONE persistent launch (Renderer.ComputeFrame x n without n launch latencies).
    public void ComputeFrames(uint n) => Check(_ctx, ptb_compute_frames(_ctx, n));
    public void ReadFloatImage(Span<float> f32) { fixed (float* p = f32) Check(_ctx, ptb_read_f32(_ctx, p, 0)); }      // needs CfgKeepF32
    public void ReadHitIds(Span<int> ids) { fixed (int* p = ids) Check(_ctx, ptb_read_hit_ids(_ctx, p, 0)); }          // needs CfgKeepHitIds
    public IntPtr DeviceFloatImage => (IntPtr)ptb_device_f32(_ctx);
    // checkpoint / restore of the progressive state (SURVEY.md 8f-4)
    public void ReadAccumulation(Span<float> acc) { fixed (float* p = acc) Check(_ctx, ptb_read_accum(_ctx, p, 0)); }
    public void WriteAccumulation(ReadOnlySpan<float> acc, uint sppTotal) { fixed (float* p = acc) Check(_ctx, ptb_write_accum(_ctx, p, 0, sppTotal)); }
    public void SetAccumBuffer(IntPtr devicePtr) => Check(_ctx, ptb_set_accum_buffer(_ctx, (void*)devicePtr));
    public void SetAccumulatedSamples(uint spp) => Check(_ctx, ptb_set_accum_spp(_ctx, spp));
    public void SetRowShardRows(uint rowOffset, uint rowStride) => Check(_ctx, ptb_set_row_shard(_ctx, rowOffset, rowStride));
    public SceneInfo GetSceneInfo() { SceneInfo i; Check(_ctx, ptb_scene_info_get(_ctx, &i)); return i; }
    public SceneStats GetSceneStats() { SceneStats s; Check(_ctx, ptb_scene_stats_get(_ctx, &s)); return s; }
    public void ResetSceneStats() => Check(_ctx, ptb_scene_stats_reset(_ctx));
    public uint[] SceneQueueSizes()
    {
        var q = new uint[256];
        uint n;
        fixed (uint* p = q) Check(_ctx, ptb_scene_queue_sizes(_ctx, p, (uint)q.Length, out n));
        Array.Resize(ref q, (int)n);
        return q;
    }
    public (float[] nodes, uint[] order) ReadBvh(int which)                // 64 B nodes as 16 floats each, leaf-order primitive ids
    {
        Check(_ctx, ptb_scene_read_bvh(_ctx, which, null, 0, out uint nn, null, 0, out uint np));
        var nodes = new float[nn * 16];
        var order = new uint[np];
        fixed (float* pn = nodes) fixed (uint* po = order)
            Check(_ctx, ptb_scene_read_bvh(_ctx, which, pn, nn, out nn, po, np, out np));
        return (nodes, order);
    }
    // cross-GPU rendezvous + the native multi-GPU frame loop
    public IntPtr DevicePeerFlags => (IntPtr)ptb_device_peer_flags(_ctx);
    public void PeerBarrier(ReadOnlySpan<IntPtr> peerFlags, uint selfRank, uint phase, uint epoch)
    {
        fixed (IntPtr* p = peerFlags) Check(_ctx, ptb_peer_barrier(_ctx, (uint)peerFlags.Length, p, selfRank, phase, epoch));
    }
    public void PeerGroupCreate() => Check(_ctx, ptb_peer_group_create(_ctx));
    public IntPtr PeerGroupLocal(int which) => (IntPtr)ptb_peer_group_local(_ctx, which);
    public void PeerGroupBind(uint selfRank, uint dstRank, ReadOnlySpan<IntPtr> partial0, ReadOnlySpan<IntPtr> partial1,
                              ReadOnlySpan<IntPtr> flags, IntPtr dstImage0, IntPtr dstImage1)
    {
        fixed (IntPtr* a = partial0, b = partial1, f = flags)
            Check(_ctx, ptb_peer_group_bind(_ctx, (uint)partial0.Length, selfRank, dstRank, a, b, f, dstImage0, dstImage1));
    }
    /// This rank's samples of one frame + the pipelined fused peer merge: ONE call per frame and rank.
    public void PeerFrame(uint firstSample, uint nSamples, uint sppTotal, bool jitter = true, uint flags = 0) =>
        Check(_ctx, ptb_peer_frame(_ctx, firstSample, nSamples, jitter ? 1 : 0, sppTotal, flags));
    public void PeerFinish() => Check(_ctx, ptb_peer_finish(_ctx));
    public float[] PeerReadTotal(out uint slabFirst)
    {
        Check(_ctx, ptb_peer_read_total(_ctx, null, out slabFirst, out uint count));
        var t = new float[count * 4];
        fixed (float* p = t) Check(_ctx, ptb_peer_read_total(_ctx, p, out slabFirst, out count));
        return t;
    }
    public void TimerBegin() => Check(_ctx, ptb_timer_begin(_ctx));
    public float TimerEnd() { Check(_ctx, ptb_timer_end(_ctx, out float ms)); return ms; }
    public (ulong mismatches, ulong tested) SelfTestExactMath(int which, float divisor = 1f)
    {
        Check(_ctx, ptb_selftest_exact_math(_ctx, which, divisor, out ulong bad, out ulong n));
        return (bad, n);
    }

    public void Dispose()
    {
        if (_ctx != IntPtr.Zero) { ptb_destroy(_ctx); _ctx = IntPtr.Zero; }
    }
}
