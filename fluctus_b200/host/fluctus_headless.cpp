// fluctus_headless — headless render entry point of the B200 backend.
//
// Replaces the reference's `Fluctus -b` batch mode (reference src/main.cpp:20-140 + Tracer::renderSingle,
// src/tracer.cpp:108-182), which cannot run without a GLFW window and an OpenCL-GL context.  The command line
// keeps the reference's flags (-x -y -s -b scenes...) and its settings.json surface.  --renderer wavefront (default)
// renders with the wavefront loop, the path this backend accelerates; --renderer mk runs Tracer::renderSingle as
// the reference's batch mode does: the microkernel renderer, exactly SPP samples in every pixel, deterministic.
//
//   fluctus_headless [-x W] [-y H] [-s SPP] [-b] [--settings settings.json] [--env map.hdr] [--bounces N]
//                    [--renderer wavefront|mk] [--tasks N] [--lbvh MAXLEAF] [--iterations N] [--device D] [--gpus G]
//                    [--rank R --world W --comm-file F]
//                    [--benchmark SECONDS [--csv report.csv]] [--benchmark-file F.bm.json]
//                    [--load-state F.dat] [--save-state F.dat]
//                    [--load-accum F.flxa] [--save-accum F.flxa]
//                    [--out image.ppm|.pfm] [--save-scene out.flxs] scene...
//
// --gpus G      sample-sharded render on devices D..D+G-1 from this one process (one HeadlessTracer and one host
//               thread per GPU, seeds offset per rank, spp split); the accumulation buffers are summed over NVLink
//               peer memory (flx_reduce_frames) and rank 0 writes the image.
// --rank R --world W --comm-file F   the same sharding with ONE PROCESS PER GPU: start W copies of this command with
//               R = 0..W-1 (device R unless --device is given).  Rank 0 draws an NCCL id and publishes it in file F, the
//               others wait for it; every rank renders its share and the accumulation buffers are summed with one
//               ncclReduce to rank 0 on the render stream (flx_comm_init / flx_reduce_pixels), which writes the image.
//               Every rank prints its own JSON line (its own ray counts); rank 0's names the image.
// --benchmark S Tracer::runBenchmark's timed loop (src/tracer.cpp:391-565) for S seconds per scene (the reference:
//               30 s at 1024x1024): CSV rows `scene;time;primary;extension;shadow;total;samples` every half second
//               to --csv, and the reference's one-line summary per scene on stdout.
// --benchmark-file F.bm.json   Tracer::runBenchmarkFromFile (src/tracer.cpp:567-870): the reference's benchmark description
//               — {"settings": {...}, "outputFolder": "...", "scenes": [{"file": ..., "outputFile": ..., "settings": {...}}]},
//               paths relative to the file — parsed by the reference's own Settings::import (built with the reference
//               tree only).  Per scene: window size, renderer (useWavefront), bounces, lights, camera ... from the
//               settings, render until maxRenderTime / maxSpp (30 s when neither is given), then
//               <outputFolder>/<outputFile>.ppm + .pfm (the reference: .png + .hdr), .csv and .txt as the reference writes
//               them, the summary line and one JSON line on stdout.  Takes no scene arguments.
// --dump-settings  print the settings in force (after --settings / per scene of --benchmark-file) as JSON and exit: no GPU work.
// --load-state / --save-state  the reference's data/states/state_<hash>.dat (camera, lights, sampling, tonemap).
// --load-accum / --save-accum  raw accumulation checkpoint: a progressive render resumes where it stopped.
//
// scene: a .flxs container (host_scene.hpp) or — when built against the reference tree — any file the
// reference's Scene::loadModel understands (.obj .ply .pbrt .pbf .sc.json), loaded and SBVH-built by the
// reference's own code.  Prints one JSON line per scene with the statistics the reference logs
// (src/tracer.cpp:184-200, :441-444): primary / extension / shadow rays, samples, seconds, MRays/s.
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <thread>
#include <unistd.h>
#include "headless_tracer.hpp"

// rank 0 publishes the NCCL id in `file` (written under a temporary name, then renamed: readers never see half a file);
// the other ranks poll for it
static std::vector<unsigned char> exchangeCommId(const std::string &file, int rank, double timeoutSeconds = 120.0)
{
    if (rank == 0)
    {
        const std::vector<unsigned char> id = flx::CudaContext::commUniqueId();
        const std::string tmp = file + ".tmp";
        { std::ofstream f(tmp, std::ios::binary); f.write((const char *)id.data(), (std::streamsize)id.size()); if (!f) throw std::runtime_error("cannot write " + tmp); }
        if (std::rename(tmp.c_str(), file.c_str()) != 0) throw std::runtime_error("cannot rename " + tmp + " to " + file);
        return id;
    }
    const auto t0 = std::chrono::steady_clock::now();
    for (;;)
    {
        std::ifstream f(file, std::ios::binary);
        std::vector<unsigned char> id(FLX_COMM_ID_BYTES);
        if (f && f.read((char *)id.data(), (std::streamsize)id.size())) return id;
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeoutSeconds)
            throw std::runtime_error("timed out waiting for rank 0's communicator id in " + file);
        usleep(20000);
    }
}

static bool endsWith(const std::string &s, const std::string &e) { return s.size() >= e.size() && s.compare(s.size() - e.size(), e.size(), e) == 0; }

static void usage()
{
    std::cerr << "usage: fluctus_headless [-x W] [-y H] [-s SPP] [-b] [--settings FILE] [--env FILE.hdr] [--bounces N]\n"
                 "                        [--renderer wavefront|mk] [--tasks N] [--lbvh MAXLEAF] [--iterations N] [--device D] [--gpus G]\n"
                 "                        [--rank R --world W --comm-file F]\n"
                 "                        [--benchmark SECONDS [--csv FILE.csv]] [--benchmark-file F.bm.json]\n"
                 "                        [--load-state F.dat] [--save-state F.dat]\n"
                 "                        [--load-accum F.flxa] [--save-accum F.flxa]\n"
                 "                        [--out FILE.ppm|.pfm] [--save-scene FILE.flxs] scene...\n";
}


// --dump-settings: the settings in force as one JSON line (what the reference's Settings getters return after the import)
static void dumpSettings(const flx::Settings &s, const std::string &scene, const std::string &output)
{
    const flx::Settings::Camera &c = s.camera; const flx::Settings::AreaLight &l = s.areaLight;
    printf("{\"scene\": \"%s\", \"output\": \"%s\", \"envMap\": \"%s\", \"renderScale\": %g, \"windowWidth\": %d, \"windowHeight\": %d, \"wfBufferSize\": %u, "
           "\"useWavefront\": %s, \"useRussianRoulette\": %s, \"useSeparateQueues\": %s, \"clUseBitstack\": %s, \"maxPathDepth\": %d, \"maxSpp\": %u, "
           "\"maxRenderTime\": %u, \"sampleImplicit\": %s, \"sampleExplicit\": %s, \"useEnvMap\": %s, \"useAreaLight\": %s, \"tonemap\": %d, "
           "\"camera\": {\"pos\": [%.9g, %.9g, %.9g], \"dir\": [%.9g, %.9g, %.9g], \"up\": [%.9g, %.9g, %.9g], \"right\": [%.9g, %.9g, %.9g], \"fov\": %.9g, "
           "\"apertureSize\": %.9g, \"focalDist\": %.9g}, \"areaLight\": {\"pos\": [%.9g, %.9g, %.9g], \"N\": [%.9g, %.9g, %.9g], \"up\": [%.9g, %.9g, %.9g], "
           "\"right\": [%.9g, %.9g, %.9g], \"E\": [%.9g, %.9g, %.9g], \"size\": [%.9g, %.9g]}, \"settings_by\": \"%s\"}\n",
           scene.c_str(), output.c_str(), s.envMapName.c_str(), s.renderScale, s.windowWidth, s.windowHeight, s.wfBufferSize,
           s.useWavefront ? "true" : "false", s.useRussianRoulette ? "true" : "false", s.useSeparateQueues ? "true" : "false", s.clUseBitstack ? "true" : "false",
           s.maxPathDepth, s.maxSpp, s.maxRenderTime, s.sampleImplicit ? "true" : "false", s.sampleExplicit ? "true" : "false",
           s.useEnvMap ? "true" : "false", s.useAreaLight ? "true" : "false", s.tonemap,
           c.pos[0], c.pos[1], c.pos[2], c.dir[0], c.dir[1], c.dir[2], c.up[0], c.up[1], c.up[2], c.right[0], c.right[1], c.right[2], c.fov, c.apertureSize, c.focalDist,
           l.pos[0], l.pos[1], l.pos[2], l.N[0], l.N[1], l.N[2], l.up[0], l.up[1], l.up[2], l.right[0], l.right[1], l.right[2], l.E[0], l.E[1], l.E[2], l.size[0], l.size[1],
           s.parsedByReference ? "reference settings.cpp" : "scanner");
}

// Tracer::runBenchmarkFromFile (src/tracer.cpp:567-870) without the window
static int runBenchmarkFile(const std::string &benchmarkFile, int device, uint32_t tasks, uint32_t lbvhLeaf, bool dumpOnly)
{
#ifndef FLX_WITH_REFERENCE_HOST
    (void)device; (void)tasks; (void)lbvhLeaf; (void)dumpOnly;
    std::cerr << "fluctus_headless: --benchmark-file needs the build with the reference tree (its Settings::import and scene loaders); got " << benchmarkFile << "\n";
    return 2;
#else
    std::vector<flx::BenchmarkScene> jobs;
    try { jobs = flx::loadBenchmarkFileWithReference(benchmarkFile); }
    catch (const std::exception &e) { std::cerr << "fluctus_headless: " << e.what() << "\n"; return 2; }
    int rc = 0;
    for (size_t k = 0; k < jobs.size(); k++)
    {
        const flx::BenchmarkScene &job = jobs[k];
        if (dumpOnly) { dumpSettings(job.settings, job.file, job.outputPrefix); continue; }
        try
        {
            flx::Settings st = job.settings;
            if (tasks) st.wfBufferSize = tasks;
            const int width = st.windowWidth, height = st.windowHeight; // resetParams(settings.getWindowWidth(), ...) at renderScale 1
            if (!st.useWavefront && (uint64_t)st.wfBufferSize < (uint64_t)width * height) st.wfBufferSize = (uint32_t)(width * height);
            flx::HostScene scene = endsWith(job.file, ".flxs") ? flx::loadScene(job.file) : flx::loadSceneWithReference(job.file, st.envMapName);
            if (lbvhLeaf) scene.nodes = flx::CudaContext::buildHierarchy(scene.tris.data(), scene.tris.size(), scene.indices, lbvhLeaf, device, nullptr);
            flx::HeadlessTracer tracer(width, height, st, device);
            tracer.init(scene);
            if (!scene.hasEnvMap()) tracer.params.useEnvMap = 0;
            std::ostringstream csv;
            const std::string summary = tracer.runBenchmarkScene((double)st.maxRenderTime, st.useWavefront, csv);
            std::cout << summary << std::endl;
            tracer.saveImage(job.outputPrefix + ".ppm");
            tracer.saveImage(job.outputPrefix + ".pfm");
            { std::ofstream f(job.outputPrefix + ".csv"); f << csv.str(); if (!f) std::cout << "Failed to write CSV benchmark report!" << std::endl; }
            { std::ofstream f(job.outputPrefix + ".txt"); f << summary << "\n"; if (!f) std::cout << "Failed to write TXT benchmark report!" << std::endl; }
            const flx::CudaContext::Totals &t = tracer.lastTotals;
            const double sec = tracer.lastSeconds;
            printf("{\"scene\": \"%s\", \"output\": \"%s\", \"width\": %d, \"height\": %d, \"triangles\": %zu, \"renderer\": \"%s\", \"bounces\": %d, "
                   "\"max_spp\": %u, \"max_render_time\": %u, \"primary\": %llu, \"extension\": %llu, \"shadow\": %llu, \"samples\": %llu, \"splatted\": %llu, "
                   "\"seconds\": %.4f, \"msamples_per_s\": %.3f, \"mrays_per_s\": %.3f, \"settings_by\": \"%s\", \"kernel_launches\": %llu}\n",
                   job.file.c_str(), job.outputPrefix.c_str(), width, height, scene.tris.size(), st.useWavefront ? "wavefront" : "mk", st.maxPathDepth,
                   st.maxSpp, st.maxRenderTime, (unsigned long long)t.primary, (unsigned long long)t.extension, (unsigned long long)t.shadow,
                   (unsigned long long)t.samples, (unsigned long long)t.splatted, sec, t.samples / std::max(sec, 1e-9) * 1e-6,
                   (t.primary + t.extension + t.shadow) / std::max(sec, 1e-9) * 1e-6, st.parsedByReference ? "reference settings.cpp" : "scanner",
                   (unsigned long long)tracer.getClContext().launchCount());
        }
        catch (const std::exception &e)
        {
            std::cerr << "fluctus_headless: " << job.file << ": " << e.what() << "\n";
            rc = 1;
        }
    }
    return rc;
#endif
}

int main(int argc, char **argv)
{
    flx::Settings settings;
    int width = -1, height = -1, spp = 32, device = -1, bounces = -1, gpus = 1, rank = 0, world = 1;
    uint32_t iterations = 0, tasks = 0, lbvhLeaf = 0;
    bool useMK = false, dumpOnly = false;
    double benchSeconds = 0.0;
    std::string out = "output.ppm", saveScene, env, csvFile, loadState, saveState, loadAccum, saveAccum, commFile, benchmarkFile;
    std::vector<std::string> scenes;
    try
    {
        for (int i = 1; i < argc; i++)
        {
            const std::string a = argv[i];
            auto next = [&]() -> std::string { if (i + 1 >= argc) throw std::runtime_error("missing value for " + a); return argv[++i]; };
            if (a == "-h" || a == "--help") { usage(); return 0; }
            else if (a == "-x" || a == "--width") width = std::stoi(next());
            else if (a == "-y" || a == "--height") height = std::stoi(next());
            else if (a == "-s" || a == "--samples") spp = std::stoi(next());
            else if (a == "-b" || a == "--batch") {} // always batch: there is no window
            else if (a == "--settings") settings.load(next());
            else if (a == "--env") env = next();
            else if (a == "--bounces") bounces = std::stoi(next());
            else if (a == "--iterations") iterations = (uint32_t)std::stoul(next());
            else if (a == "--device") device = std::stoi(next());
            else if (a == "--renderer") { const std::string r = next(); if (r != "mk" && r != "wavefront") throw std::runtime_error("--renderer mk|wavefront"); useMK = (r == "mk"); }
            else if (a == "--tasks") tasks = (uint32_t)std::stoul(next());
            else if (a == "--lbvh") lbvhLeaf = (uint32_t)std::stoul(next());
            else if (a == "--gpus") gpus = std::stoi(next());
            else if (a == "--rank") rank = std::stoi(next());
            else if (a == "--world") world = std::stoi(next());
            else if (a == "--comm-file") commFile = next();
            else if (a == "--benchmark") benchSeconds = std::stod(next());
            else if (a == "--csv") csvFile = next();
            else if (a == "--benchmark-file") benchmarkFile = next();
            else if (a == "--dump-settings") dumpOnly = true;
            else if (a == "--load-state") loadState = next();
            else if (a == "--save-state") saveState = next();
            else if (a == "--load-accum") loadAccum = next();
            else if (a == "--save-accum") saveAccum = next();
            else if (a == "--out") out = next();
            else if (a == "--save-scene") saveScene = next();
            else if (!a.empty() && a[0] == '-') throw std::runtime_error("unknown option " + a);
            else scenes.push_back(a);
        }
        if (!benchmarkFile.empty())
        {
            if (!scenes.empty() || gpus > 1 || world > 1 || benchSeconds > 0.0 || !loadAccum.empty())
                throw std::runtime_error("--benchmark-file names its own scenes and runs on one GPU (no scene arguments, --gpus, --world, --benchmark, --load-accum)");
        }
        else if (scenes.empty() && !dumpOnly) throw std::runtime_error("no scene given");
        if (width < 0) width = settings.windowWidth;
        if (height < 0) height = settings.windowHeight;
        if (width <= 0 || height <= 0 || spp < 0) throw std::runtime_error("invalid width / height / samples");
        if (gpus < 1 || gpus > 16) throw std::runtime_error("--gpus 1..16");
        if (benchSeconds < 0.0) throw std::runtime_error("--benchmark needs a positive duration");
        if (gpus > 1 && (benchSeconds > 0.0 || !loadAccum.empty())) throw std::runtime_error("--benchmark / --load-accum run on one GPU");
        if (world < 1 || world > 64 || rank < 0 || rank >= world) throw std::runtime_error("--rank R --world W with 0 <= R < W <= 64");
        if (world > 1 && commFile.empty()) throw std::runtime_error("--world needs --comm-file (where rank 0 publishes the NCCL id)");
        if (world > 1 && (gpus > 1 || benchSeconds > 0.0 || !loadAccum.empty()))
            throw std::runtime_error("--world (one process per GPU) excludes --gpus / --benchmark / --load-accum");
        if (device < 0) device = world > 1 ? rank : 0;
        if (bounces >= 0) settings.maxPathDepth = bounces;
        if (env.empty()) env = settings.envMapName;
    }
    catch (const std::exception &e)
    {
        std::cerr << "fluctus_headless: " << e.what() << "\n";
        usage();
        return 2;
    }

    if (!benchmarkFile.empty()) return runBenchmarkFile(benchmarkFile, device, tasks, lbvhLeaf, dumpOnly);
    if (dumpOnly) { dumpSettings(settings, scenes.empty() ? "" : scenes[0], out); return 0; }

    int rc = 0;
    for (size_t si = 0; si < scenes.size(); si++)
    {
        try
        {
            flx::HostScene scene;
            if (endsWith(scenes[si], ".flxs")) scene = flx::loadScene(scenes[si]);
            else
            {
#ifdef FLX_WITH_REFERENCE_HOST
                scene = flx::loadSceneWithReference(scenes[si], env);
#else
                throw std::runtime_error("this build reads .flxs scenes only (built without the reference tree); got " + scenes[si]);
#endif
            }
            if (lbvhLeaf) // replace the hierarchy that came with the scene by one built on the GPU (flx_build_lbvh)
            {
                uint32_t depth = 0;
                scene.nodes = flx::CudaContext::buildHierarchy(scene.tris.data(), scene.tris.size(), scene.indices, lbvhLeaf, device, &depth);
                std::cerr << "rebuilt hierarchy on the GPU: " << scene.nodes.size() << " nodes, depth " << depth << "\n";
            }
            if (!saveScene.empty()) flx::saveScene(saveScene, scene);

            // Tracer::renderSingle disables Russian roulette for fixed-spp renders (src/tracer.cpp:117-121)
            flx::Settings sceneSettings = settings;
            if (tasks) sceneSettings.wfBufferSize = tasks;
            // the microkernel renderer keeps one task per pixel (mk_raygen.cl:8-9)
            if (useMK && (uint64_t)sceneSettings.wfBufferSize < (uint64_t)width * height) sceneSettings.wfBufferSize = (uint32_t)(width * height);
            // one tracer per GPU: rank r renders the samples of shard r (seed offset r*N, pixel cursor r*w*h/G, spp/G)
            // (--gpus: the shards are this process's tracers; --world: this process is shard `rank` of `world`)
            const int shards = world > 1 ? world : gpus;
            auto shardOf = [&](int r) { return world > 1 ? rank : r; };
            if (shards > 1 && sceneSettings.maxSpp > 0) sceneSettings.maxSpp = (sceneSettings.maxSpp + shards - 1) / shards;
            std::vector<std::unique_ptr<flx::HeadlessTracer>> tracers;
            for (int r = 0; r < gpus; r++) tracers.emplace_back(new flx::HeadlessTracer(width, height, sceneSettings, device + r, shardOf(r), shards));
            flx::HeadlessTracer &tracer = *tracers[0];
            for (auto &t : tracers) { t->init(scene); if (!loadState.empty()) t->loadState(loadState); }
            if (!saveState.empty()) tracer.saveState(saveState);
            if (gpus > 1) // map the peers' frame buffers now (a combine of the still empty frames): a one-off ~40 ms per GPU pair
            {
                std::vector<flx::CudaContext *> ctxs;
                for (auto &tr : tracers) ctxs.push_back(&tr->getClContext());
                flx::CudaContext::reduceFrames(ctxs);
            }
            if (world > 1) // join the communicator; one combine of the still empty frames sets NCCL's connections up and lines the ranks up
            {
                flx::CudaContext &c = tracer.getClContext();
                c.commInit(rank, world, exchangeCommId(commFile + (scenes.size() > 1 ? "." + std::to_string(si) : ""), rank));
                c.reducePixels(0); // the collective used below: NCCL sets its connections up per algorithm at the first call (1.2 s at 8 ranks)
                c.finishQueue();
            }
            const auto t0 = std::chrono::steady_clock::now();
            flx::CudaContext::Totals t{};
            if (benchSeconds > 0.0) // Tracer::runBenchmark
            {
                std::ofstream csv;
                if (!csvFile.empty())
                {
                    csv.open(csvFile, si == 0 ? std::ios::out : std::ios::app);
                    if (!csv) throw std::runtime_error("cannot write " + csvFile);
                    if (si == 0) csv << "scene;time;primary;extension;shadow;total;samples\n";
                }
                std::ostringstream sink;
                std::cout << tracer.runBenchmark(scenes[si], benchSeconds, csvFile.empty() ? (std::ostream &)sink : (std::ostream &)csv) << std::endl;
                t = tracer.getClContext().getRenderTotals();
            }
            else
            {
                std::vector<flx::CudaContext::Totals> per(gpus);
                std::vector<std::string> errors(gpus);
                auto work = [&](int r)
                {
                    try
                    {
                        flx::HeadlessTracer &tr = *tracers[r];
                        if (useMK) // Tracer::renderSingle — what the reference's own `-b -s SPP` runs (src/main.cpp, tracer.cpp:108-182)
                        {
                            const flx::RenderStats st = tr.renderSingle((uint32_t)spp); // takes this rank's share of spp itself
                            per[r] = flx::CudaContext::Totals{st.primaryRays, st.extensionRays, st.shadowRays, st.samples, 0, (uint64_t)spp};
                        }
                        else
                        {
                            const uint32_t ws = (uint32_t)shards, me = (uint32_t)shardOf(r);
                            const uint32_t share = (uint32_t)spp / ws + (me < (uint32_t)spp % ws ? 1u : 0u);
                            if (!loadAccum.empty())
                            {
                                tr.render(0, 0); // reset + priming segment, then the checkpoint goes on top
                                tr.loadAccumulation(loadAccum);
                                per[r] = tr.resume(share, iterations ? iterations : (1u << 20));
                            }
                            else per[r] = tr.render(share, iterations ? iterations : (1u << 20));
                        }
                    }
                    catch (const std::exception &e) { errors[r] = e.what(); }
                };
                if (gpus == 1) work(0);
                else
                {
                    std::vector<std::thread> threads;
                    for (int r = 0; r < gpus; r++) threads.emplace_back(work, r);
                    for (auto &th : threads) th.join();
                }
                for (int r = 0; r < gpus; r++) if (!errors[r].empty()) throw std::runtime_error("GPU " + std::to_string(device + r) + ": " + errors[r]);
                for (int r = 0; r < gpus; r++)
                {
                    t.primary += per[r].primary; t.extension += per[r].extension; t.shadow += per[r].shadow; t.samples += per[r].samples;
                    t.splatted += per[r].splatted; t.iterations = std::max(t.iterations, per[r].iterations);
                }
                if (gpus > 1) // the frame = sum of the shards' accumulation buffers, left in rank 0's context
                {
                    std::vector<flx::CudaContext *> ctxs;
                    for (auto &tr : tracers) ctxs.push_back(&tr->getClContext());
                    flx::CudaContext::reduceFrames(ctxs);
                }
                if (world > 1) // one ncclReduce on the render stream: rank 0's buffer becomes the frame
                {
                    tracer.getClContext().reducePixels(0);
                    tracer.getClContext().finishQueue();
                }
            }
            if (!saveAccum.empty() && rank == 0) tracer.saveAccumulation(saveAccum);
            const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            std::string file = out;
            if (scenes.size() > 1) file = out.substr(0, out.rfind('.')) + "_" + std::to_string(si) + out.substr(out.rfind('.'));
            if (rank == 0) tracer.saveImage(file);
            uint64_t launches = 0;
            for (auto &tr : tracers) launches += tr->getClContext().launchCount();
            printf("{\"scene\": \"%s\", \"width\": %d, \"height\": %d, \"triangles\": %zu, \"iterations\": %llu, \"primary\": %llu, "
                   "\"extension\": %llu, \"shadow\": %llu, \"samples\": %llu, \"seconds\": %.4f, \"msamples_per_s\": %.3f, "
                   "\"mrays_per_s\": %.3f, \"gpus\": %d, \"rank\": %d, \"world\": %d, \"nccl\": %d, \"kernel_launches\": %llu, \"image\": \"%s\"}\n",
                   scenes[si].c_str(), width, height, scene.tris.size(), (unsigned long long)t.iterations, (unsigned long long)t.primary,
                   (unsigned long long)t.extension, (unsigned long long)t.shadow, (unsigned long long)t.samples, sec, t.samples / sec * 1e-6,
                   (t.primary + t.extension + t.shadow) / sec * 1e-6, gpus, rank, world, world > 1 ? tracer.getClContext().ncclVersion() : 0,
                   (unsigned long long)launches, rank == 0 ? file.c_str() : "");
        }
        catch (const std::exception &e)
        {
            // the reference's loaders call exit(EXIT_FAILURE) (src/utils.h:28-34); CLContext errors throw runtime_error
            std::cerr << "fluctus_headless: " << scenes[si] << ": " << e.what() << "\n";
            rc = 1;
        }
    }
    return rc;
}
