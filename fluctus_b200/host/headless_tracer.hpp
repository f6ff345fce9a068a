// HeadlessTracer — the reference's render driver (`class Tracer`, reference src/tracer.{hpp,cpp}) without the
// GLFW window: same settings surface (settings.json keys of src/settings.cpp:89-128 with the defaults of
// :17-59), same parameter initialisation (resetParams :43-65, initCamera / initPostProcessing / initAreaLight
// :1127-1168), same stage sequence (update :250-294 + :333-368; benchmark loop :633-645, :725-778), same image
// export arithmetic (CLContext::saveImage, src/clcontext.cpp:404-483).  The device side is CudaContext.
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <fstream>
#include <sstream>
#include <string>
#include "cuda_context.hpp"
#include "host_scene.hpp"
#include "host_settings.hpp"

namespace flx {

// cam.fovSCALE = tan(toRad(0.5f * cam.fov)) (src/tracer.cpp:1137): the argument in float as there, the tangent in double
// and rounded once — independent of which libm tanf a host has, and what the Python mirror computes (abi.py)
inline float fovScale(float fov) { return (float)std::tan((double)(0.5f * fov * 0.01745329251994329576923690768489f)); }
inline flx_float3 f3(float x, float y, float z, float w = 0.0f) { flx_float3 v = {x, y, z, w}; return v; }

class HeadlessTracer
{
public:
    RenderParams params{};
    Settings settings;
    uint32_t iteration = 0;

    HeadlessTracer(int width, int height, const Settings &s = Settings(), int device = 0, int rank = 0, int worldSize = 1)
        : settings(s), clctx(s.wfBufferSize, device), rank(rank), worldSize(worldSize)
    {
        resetParams(width, height);
        initCamera();
        initPostProcessing();
        initAreaLight();
    }

    CudaContext &getClContext() { return clctx; }

    // Tracer::resetParams (src/tracer.cpp:43-65)
    void resetParams(int width, int height)
    {
        params.width = (uint32_t)(width * settings.renderScale);
        params.height = (uint32_t)(height * settings.renderScale);
        params.useEnvMap = settings.useEnvMap; params.useAreaLight = settings.useAreaLight;
        params.envMapStrength = 1.0f;
        params.maxBounces = (uint32_t)settings.maxPathDepth;
        params.sampleImpl = settings.sampleImplicit; params.sampleExpl = settings.sampleExplicit;
        params.useRoulette = settings.useRussianRoulette; params.wfSeparateQueues = settings.useSeparateQueues;
        params.maxSpp = settings.maxSpp;
        params.width1 = 1.0f / (float)params.width; params.height1 = 1.0f / (float)params.height;
    }
    // Tracer::initCamera / initPostProcessing / initAreaLight (src/tracer.cpp:1127-1168), defaults src/settings.cpp:40-58
    void initCamera()
    {
        flx_camera &c = params.camera;
        const Settings::Camera &s = settings.camera; // Settings::getCameraSettings()
        c.pos = f3(s.pos[0], s.pos[1], s.pos[2]); c.right = f3(s.right[0], s.right[1], s.right[2]);
        c.up = f3(s.up[0], s.up[1], s.up[2]); c.dir = f3(s.dir[0], s.dir[1], s.dir[2]);
        c.fov = s.fov; c.fovSCALE = fovScale(c.fov);
        c.apertureSize = s.apertureSize; c.focalDist = s.focalDist;
    }
    void setCamera(const float pos[3], const float dir[3], const float up[3], const float right[3], float fov)
    {
        flx_camera &c = params.camera;
        c.pos = f3(pos[0], pos[1], pos[2]); c.dir = f3(dir[0], dir[1], dir[2]); c.up = f3(up[0], up[1], up[2]); c.right = f3(right[0], right[1], right[2]);
        c.fov = fov; c.fovSCALE = fovScale(fov);
    }
    void initPostProcessing() { params.ppParams.exposure = 1.0f; params.ppParams.tmOperator = (uint32_t)settings.tonemap; }
    void initAreaLight()
    {
        flx_area_light &l = params.areaLight;
        const Settings::AreaLight &s = settings.areaLight; // Settings::getAreaLightSettings()
        l.right = f3(s.right[0], s.right[1], s.right[2]); l.up = f3(s.up[0], s.up[1], s.up[2]);
        l.N = f3(s.N[0], s.N[1], s.N[2], s.N[3]); l.pos = f3(s.pos[0], s.pos[1], s.pos[2], s.pos[3]);
        l.E = f3(s.E[0], s.E[1], s.E[2]); l.size.x = s.size[0]; l.size.y = s.size[1];
    }

    // Tracer::init (src/tracer.cpp:68-93): scene -> env map -> hierarchy -> worldRadius -> upload
    void init(const HostScene &s)
    {
        if (s.hasEnvMap())
        {
            clctx.createEnvMap(s.envRgb.data(), s.envW, s.envH, s.envProb.data(), s.envAlias.data(), s.envPdf.data());
            params.useEnvMap = 1; // Tracer::initEnvMap (src/tracer.cpp:922-931)
        }
        else params.useEnvMap = 0;
        params.n_tris = (uint32_t)s.tris.size();
        params.worldRadius = s.worldRadius();
        clctx.uploadSceneData(s.tris.data(), s.tris.size(), s.indices.data(), s.indices.size(), s.nodes.data(), s.nodes.size(),
                              s.materials.data(), s.materials.size(), s.textures.empty() ? nullptr : s.textures.data(), s.textures.size(),
                              s.texels.empty() ? nullptr : s.texels.data(), s.texels.size());
        clctx.setTraversal(settings.clUseBitstack); // -DUSE_BITSTACK (src/clcontext.cpp:145-155)
        clctx.setOptions(false, (uint32_t)rank * clctx.getNumTasks()); // sample sharding: seeds of one worldSize*N-task run
        clctx.updateParams(params);
        iteration = 0;
    }

    uint32_t cursorStart() const { return (uint32_t)(((uint64_t)rank * params.width * params.height) / (uint64_t)worldSize); }

    // Tracer::update, wavefront branch (src/tracer.cpp:250-294, :333-336, :361-368): one interactive iteration
    QueueCounters update()
    {
        const uint32_t maxBounces = params.maxBounces;
        int N = 1;
        if (iteration == 0)
        {
            params.maxBounces = std::min(2u, maxBounces); // 2-bounce preview
            clctx.updateParams(params);
            N = 3;
            clctx.setPixelIndex(cursorStart());
            clctx.enqueueWfResetKernel(params);
            clctx.enqueueWfRaygenKernel(params);
            clctx.enqueueWfExtRayKernel(params);
            clctx.enqueueClearWfQueues();
        }
        QueueCounters cnt{};
        for (int i = 0; i < N; i++)
        {
            clctx.enqueueWfLogicKernel(params, iteration == 0);
            clctx.enqueueWfRaygenKernel(params);
            clctx.enqueueWfMaterialKernels(params);
            clctx.enqueueGetCounters(&cnt);
            clctx.enqueueWfExtRayKernel(params);
            clctx.enqueueWfShadowRayKernel(params);
            clctx.enqueueClearWfQueues();
        }
        if (iteration == 0) { params.maxBounces = maxBounces; clctx.updateParams(params); }
        clctx.enqueuePostprocessKernel(params);
        clctx.finishQueue();
        clctx.updatePixelIndex(params.width * params.height, cnt.raygenQueue);
        clctx.statsAsync.extensionRays += cnt.extensionQueue;
        clctx.statsAsync.shadowRays += cnt.shadowQueue;
        clctx.statsAsync.primaryRays += cnt.raygenQueue;
        clctx.statsAsync.samples += (iteration > 0) ? cnt.raygenQueue : 0;
        iteration++;
        return cnt;
    }

    // Fixed-budget render: Tracer::runBenchmarkFromFile's loop (src/tracer.cpp:633-645, :725-778) with the per-
    // iteration host work moved to the device.  Stops after `spp` samples per pixel on this rank (counted like the
    // reference: splattedSamples when maxSpp > 0, else regenerated paths) or after maxIterations passes.
    CudaContext::Totals render(uint32_t spp, uint32_t maxIterations = 1u << 20, uint32_t batch = 16)
    {
        clctx.updateParams(params);
        clctx.setPixelIndex(0);
        clctx.render(0, true);
        clctx.setPixelIndex(cursorStart());
        if (maxIterations == 0) { clctx.finishQueue(); return clctx.getRenderTotals(); } // reset + priming segment only
        return resume(spp, maxIterations, batch);
    }
    // continue an already reset renderer (after render(), loadAccumulation(), or a previous resume())
    CudaContext::Totals resume(uint32_t spp, uint32_t maxIterations = 1u << 20, uint32_t batch = 16)
    {
        const uint64_t target = (uint64_t)spp * params.width * params.height;
        CudaContext::Totals t{};
        uint64_t lastSplatted = ~0ull;
        for (uint32_t done = 0; done < maxIterations; done += batch)
        {
            clctx.render(std::min(batch, maxIterations - done));
            t = clctx.getRenderTotals();
            if (spp && (params.maxSpp > 0 ? t.splatted : t.samples) >= target) break;
            // CHECK_SPP budget exhausted: every pixel is saturated, nothing was splatted in a whole batch.  The
            // reference's stop rule `sampleCount >= w*h*maxSpp` (src/tracer.cpp:713,730,773) can then never be met — its
            // first logic pass saturates samplesPerPixel[0] without splatting, so the total
            // tops out at target - maxSpp unless racing paths slip extra samples in — and the loop would spin forever.
            if (params.maxSpp > 0 && t.splatted == lastSplatted) break;
            lastSplatted = t.splatted;
        }
        clctx.enqueuePostprocessKernel(params);
        clctx.finishQueue();
        return t;
    }

    // Tracer::renderSingle (src/tracer.cpp:108-182): the reference's batch mode (`-b -s SPP`, src/main.cpp) — the
    // microkernel renderer, every pixel gets exactly `spp` samples (this rank's share of them), Russian roulette
    // off, deterministic image.  Needs wfBufferSize >= width * height (mk_raygen.cl:8-9).
    RenderStats renderSingle(uint32_t spp)
    {
        if ((uint64_t)clctx.getNumTasks() < (uint64_t)params.width * params.height)
            throw std::runtime_error("renderSingle: wfBufferSize must be >= width * height for the microkernel renderer");
        if (params.useRoulette) { std::cout << "Turning off russian roulette" << std::endl; params.useRoulette = 0; }
        clctx.updateParams(params);
        const uint32_t ws = (uint32_t)worldSize, share = spp / ws + ((uint32_t)rank < spp % ws ? 1u : 0u);
        clctx.renderSingle(share, true);
        clctx.enqueuePostprocessKernel(params);
        clctx.fetchStats();
        return clctx.statsAsync;
    }

    // Tracer::runBenchmark's timed loop (src/tracer.cpp:405-417 resetRenderer, :459-519): render for `seconds`,
    // log the statistics every half second as the reference's CSV rows
    //     scene;time;primary;extension;shadow;total;samples        (M rays or samples per second of that interval)
    // and return the whole-run line `<scene>: %.1fM primary, ...` (:537-540).  The reference polls the clock after
    // every iteration because each one ends in clFinish; here `batch` iterations are replayed from the CUDA graph
    // between two polls (one host round trip per batch, not per iteration).
    std::string runBenchmark(const std::string &sceneName, double seconds, std::ostream &csv, uint32_t batch = 8)
    {
        using clock = std::chrono::steady_clock;
        auto now = [] { return std::chrono::duration<double>(clock::now().time_since_epoch()).count(); };
        clctx.updateParams(params);
        clctx.setPixelIndex(0);
        clctx.render(0, true); // reset + priming segment
        clctx.setPixelIndex(cursorStart());
        clctx.finishQueue();
        CudaContext::Totals prev = clctx.getRenderTotals(), first = prev, t = prev;
        const double startT = now();
        double lastLogTime = startT, currT = startT;
        auto logStats = [&](double elapsed, double deltaT)
        {
            const double s = 1e6 * deltaT;
            const double pr = double(t.primary - prev.primary), ex = double(t.extension - prev.extension), sh = double(t.shadow - prev.shadow);
            csv << sceneName << ";" << elapsed << ";" << pr / s << ";" << ex / s << ";" << sh / s << ";" << (pr + ex + sh) / s << ";"
                << double(t.samples - prev.samples) / s << "\n";
            prev = t; lastLogTime = currT;
        };
        while (currT - startT < seconds)
        {
            clctx.render(batch);
            t = clctx.getRenderTotals(); // synchronises
            iteration += batch;
            currT = now();
            if (currT - lastLogTime > 0.5) logStats(currT - startT, currT - lastLogTime);
        }
        if (currT > lastLogTime) logStats(currT - startT, currT - lastLogTime);
        clctx.enqueuePostprocessKernel(params);
        clctx.finishQueue();
        const double scale = 1e6 * (currT - startT);
        const double prim = double(t.primary - first.primary) / scale, ext = double(t.extension - first.extension) / scale,
                     shdw = double(t.shadow - first.shadow) / scale, samp = double(t.samples - first.samples) / scale;
        char line[512];
        snprintf(line, sizeof(line), "%s: %.1fM primary, %.2fM extension, %.2fM shadow, %.2fM samples, total: %.2fM rays/s",
                 sceneName.c_str(), prim, ext, shdw, samp, prim + ext + shdw);
        return line;
    }

    // One scene of Tracer::runBenchmarkFromFile (src/tracer.cpp:700-870): render until maxRenderTime seconds have passed
    // or w*h*maxSpp samples are in the frame, whichever comes first (30 s when neither is set, :706-711), with the
    // renderer the settings name (useWavefront; else the microkernel one, a whole sample per pixel per step), logging
    // `time;primary;extension;shadow;total;samples` every half second (:683-697) and returning the summary line (:833).
    // Post-processing runs once at the end (the reference's per-iteration tonemap exists for its window; skipPP is moot).
    std::string runBenchmarkScene(double maxRenderTime, bool wavefront, std::ostream &csv, uint32_t batch = 8)
    {
        using clock = std::chrono::steady_clock;
        auto now = [] { return std::chrono::duration<double>(clock::now().time_since_epoch()).count(); };
        if (maxRenderTime == 0.0 && params.maxSpp == 0) maxRenderTime = 30.0;
        const uint64_t maxSampleCount = (uint64_t)params.width * params.height * params.maxSpp;
        if (!wavefront)
        {
            if ((uint64_t)clctx.getNumTasks() < (uint64_t)params.width * params.height)
                throw std::runtime_error("microkernel renderer: wfBufferSize must be >= width * height");
            params.useRoulette = 0; // the renderSingle graph has a fixed number of bounces (src/tracer.cpp:118-122)
        }
        clctx.updateParams(params);
        struct Tot { uint64_t primary, extension, shadow, samples, counted; };
        auto read = [&]() -> Tot
        {
            if (wavefront) { const CudaContext::Totals t = clctx.getRenderTotals(); return Tot{t.primary, t.extension, t.shadow, t.samples, t.splatted}; }
            clctx.fetchStats();
            const RenderStats &st = clctx.statsAsync;
            return Tot{st.primaryRays, st.extensionRays, st.shadowRays, st.samples, st.samples};
        };
        if (wavefront)
        {
            clctx.setPixelIndex(0);
            clctx.render(0, true);
            clctx.setPixelIndex(cursorStart());
        }
        else clctx.renderSingle(0, true);
        clctx.finishQueue();
        Tot first = read(), prev = first, t = first;
        csv << "time;primary;extension;shadow;total;samples\n";
        const double startT = now();
        double lastLogTime = startT, currT = startT;
        auto logStats = [&](double elapsed, double deltaT)
        {
            const double s = 1e6 * deltaT;
            const double pr = double(t.primary - prev.primary), ex = double(t.extension - prev.extension), sh = double(t.shadow - prev.shadow);
            csv << elapsed << ";" << pr / s << ";" << ex / s << ";" << sh / s << ";" << (pr + ex + sh) / s << ";" << double(t.samples - prev.samples) / s << "\n";
            prev = t; lastLogTime = currT;
        };
        uint64_t lastCounted = ~0ull;
        while ((maxRenderTime == 0.0 || currT - startT < maxRenderTime) && (maxSampleCount == 0 || t.counted - first.counted < maxSampleCount))
        {
            if (wavefront) clctx.render(batch); else clctx.renderSingle(1, false);
            t = read(); // synchronises
            iteration += wavefront ? batch : 1;
            currT = now();
            if (currT - lastLogTime > 0.5) logStats(currT - startT, currT - lastLogTime);
            if (wavefront && maxSampleCount && t.counted == lastCounted) break; // CHECK_SPP budget used up (quirk 13, see resume())
            lastCounted = t.counted;
        }
        if (currT > lastLogTime) logStats(currT - startT, currT - lastLogTime);
        clctx.enqueuePostprocessKernel(params);
        clctx.finishQueue();
        lastTotals = CudaContext::Totals{t.primary - first.primary, t.extension - first.extension, t.shadow - first.shadow, t.samples - first.samples,
                                         t.counted - first.counted, iteration};
        lastSeconds = currT - startT;
        const double scale = 1e6 * std::max(currT - startT, 1e-9);
        const double prim = lastTotals.primary / scale, ext = lastTotals.extension / scale, shdw = lastTotals.shadow / scale, samp = lastTotals.samples / scale;
        char line[512];
        snprintf(line, sizeof(line), "%.1fM primary, %.2fM extension, %.2fM shadow, %.2fM samples, total: %.2fM rays/s", prim, ext, shdw, samp, prim + ext + shdw);
        return line;
    }
    CudaContext::Totals lastTotals{}; double lastSeconds = 0.0; // of the last runBenchmarkScene

    // Tracer::saveState / loadState (src/tracer.cpp:996-1054, keys F2/F3): camera, lights, sampling switches and
    // post-processing as one flat little-endian record, field order as in iterateStateItems, so a
    // data/states/state_<hash>.dat written by the reference's viewer positions this renderer's camera and vice versa.
    // cameraRotation / cameraSpeed belong to the viewer's mouse handling: kept and written back unchanged.
    void saveState(const std::string &file) { stateIO(file, true); }
    void loadState(const std::string &file) { stateIO(file, false); clctx.updateParams(params); iteration = 0; }

    // Accumulation checkpoint (new): the raw float4 buffer carries its own sample count in .w, so a
    // progressive render resumes by loading it back before the next iterations.  Header: "FLXA", width, height.
    void saveAccumulation(const std::string &file)
    {
        const std::vector<float> px = clctx.readPixels(FLX_BUF_PIXELS);
        std::ofstream o(file, std::ios::binary);
        if (!o) throw std::runtime_error("cannot write " + file);
        const uint32_t hdr[3] = {0x41584c46u, params.width, params.height};
        o.write((const char *)hdr, sizeof(hdr));
        o.write((const char *)px.data(), (std::streamsize)(px.size() * sizeof(float)));
    }
    // call after the first render()/update() has reset the renderer (a reset clears the buffer)
    void loadAccumulation(const std::string &file)
    {
        std::ifstream in(file, std::ios::binary);
        uint32_t hdr[3] = {0, 0, 0};
        if (!in || !in.read((char *)hdr, sizeof(hdr)) || hdr[0] != 0x41584c46u) throw std::runtime_error("not an accumulation checkpoint: " + file);
        if (hdr[1] != params.width || hdr[2] != params.height) throw std::runtime_error("accumulation checkpoint has another resolution: " + file);
        std::vector<float> px((size_t)params.width * params.height * 4);
        if (!in.read((char *)px.data(), (std::streamsize)(px.size() * sizeof(float)))) throw std::runtime_error("truncated accumulation checkpoint: " + file);
        clctx.writePixels(FLX_BUF_PIXELS, px);
    }

    // CLContext::saveImage (src/clcontext.cpp:404-483).  ".ppm": tonemapped preview, clamp, (uchar)(255*x), flipped
    // vertically (the reference's image origin is bottom-left).  ".pfm": raw accumulation / sample count (the
    // reference's ".hdr" branch), float RGB, bottom-up as PFM defines it.
    void saveImage(const std::string &filename)
    {
        const uint32_t w = params.width, h = params.height;
        const bool pfm = filename.size() > 4 && filename.substr(filename.size() - 4) == ".pfm";
        FILE *f = fopen(filename.c_str(), "wb");
        if (!f) throw std::runtime_error("cannot write " + filename);
        if (pfm)
        {
            const std::vector<float> px = clctx.readPixels(FLX_BUF_PIXELS);
            fprintf(f, "PF\n%u %u\n-1.0\n", w, h);
            std::vector<float> row(3 * (size_t)w);
            for (uint32_t y = 0; y < h; y++)
            {
                for (uint32_t x = 0; x < w; x++)
                {
                    const float *p = &px[4 * ((size_t)y * w + x)];
                    const float a = p[3] > 0.0f ? p[3] : 1.0f;
                    row[3 * x] = p[0] / a; row[3 * x + 1] = p[1] / a; row[3 * x + 2] = p[2] / a;
                }
                fwrite(row.data(), 4, row.size(), f);
            }
        }
        else
        {
            clctx.enqueuePostprocessKernel(params);
            const std::vector<float> px = clctx.readPixels(FLX_BUF_PREVIEW);
            fprintf(f, "P6\n%u %u\n255\n", w, h);
            std::vector<unsigned char> row(3 * (size_t)w);
            for (uint32_t y = 0; y < h; y++)
            {
                const uint32_t sy = h - 1 - y;
                for (uint32_t x = 0; x < w; x++)
                    for (int c = 0; c < 3; c++)
                    {
                        float v = px[4 * ((size_t)sy * w + x) + c];
                        v = v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v);
                        row[3 * x + c] = (unsigned char)(255.0f * v);
                    }
                fwrite(row.data(), 1, row.size(), f);
            }
        }
        fclose(f);
    }

    float cameraRotationX = 0.0f, cameraRotationY = 0.0f, cameraSpeed = 1.0f; // viewer state carried through state files

private:
    void stateIO(const std::string &file, bool write)
    {
        std::fstream io(file, std::ios::binary | (write ? std::ios::out : std::ios::in));
        if (!io.good()) throw std::runtime_error("could not open state file " + file);
        auto rw = [&](void *p, size_t n) { if (write) io.write((const char *)p, (std::streamsize)n); else io.read((char *)p, (std::streamsize)n); };
        auto f = [&](float &v) { rw(&v, 4); };
        auto u = [&](uint32_t &v) { rw(&v, 4); };
        auto vec = [&](flx_float3 &v) { rw(&v.x, 4); rw(&v.y, 4); rw(&v.z, 4); };
        flx_camera &c = params.camera; flx_area_light &l = params.areaLight;
        f(cameraRotationX); f(cameraRotationY); f(cameraSpeed); f(c.fov);
        if (!write) c.fovSCALE = fovScale(c.fov);
        f(c.focalDist); f(c.apertureSize); vec(c.dir); vec(c.pos); vec(c.right); vec(c.up);
        vec(l.N); vec(l.pos); vec(l.right); vec(l.up); vec(l.E); f(l.size.x); f(l.size.y); f(params.envMapStrength);
        u(params.maxBounces); u(params.useAreaLight); u(params.useEnvMap); u(params.sampleExpl); u(params.sampleImpl); u(params.useRoulette);
        f(params.ppParams.exposure); u(params.ppParams.tmOperator);
        if (!io.good()) throw std::runtime_error("state file " + file + (write ? ": write failed" : ": truncated"));
    }
    CudaContext clctx; // name kept from the reference (src/tracer.hpp:131)
    int rank, worldSize;
};

} // namespace flx
