// settings.json / .bm.json through the reference's OWN Settings class (reference src/settings.cpp, compiled in place
// from $FLUCTUS_REF — see Makefile; nothing is copied): Settings::import parses every key the reference knows, this file
// only drives it and copies the getters' values into the plain flx::Settings the headless host uses.
//
//   loadSettingsWithReference      Settings::load (src/settings.cpp:61-87) for a file named on the command line: the
//                                  "release" object is imported ("debug" on top in _DEBUG builds, as there)
//   loadBenchmarkFileWithReference the input half of Tracer::runBenchmarkFromFile (src/tracer.cpp:567-700): base
//                                  "settings", "outputFolder", and per scene "file" / "outputFile" / "settings"; envMap
//                                  paths relative to the benchmark file; per scene the base settings are imported again
//                                  and the scene's on top, into the same singleton (so, as in the reference, a key that
//                                  only an earlier scene set stays set)
//
// The singleton's constructor reads ./settings.json, as the reference's executable does on start-up, and prints to
// stdout when there is none; stdout carries this binary's JSON result lines, so it is silenced for that moment.
#include <fstream>
#include <iostream>
#include <sstream>
#include "settings.hpp" // the reference's
#include "utils.h"
#define FLX_WITH_REFERENCE_HOST 1
#include "host_settings.hpp"

using nlohmann::json;

namespace {

::Settings &referenceSettings()
{
    static bool first = true;
    if (first)
    {
        first = false;
        std::ostringstream sink;
        std::streambuf *old = std::cout.rdbuf(sink.rdbuf());
        ::Settings::getInstance();
        std::cout.rdbuf(old);
    }
    return ::Settings::getInstance();
}

void snapshot(flx::Settings &o)
{
    ::Settings &s = referenceSettings();
    o.envMapName = s.getEnvMapName();
    o.renderScale = s.getRenderScale();
    o.windowWidth = s.getWindowWidth(); o.windowHeight = s.getWindowHeight();
    o.wfBufferSize = s.getWfBufferSize();
    o.useWavefront = s.getUseWavefront();
    o.useRussianRoulette = s.getUseRussianRoulette(); o.useSeparateQueues = s.getUseSeparateQueues(); o.clUseBitstack = s.getUseBitstack();
    o.maxPathDepth = s.getMaxPathDepth();
    o.maxSpp = s.getMaxSpp(); o.maxRenderTime = s.getMaxRenderTime();
    o.sampleImplicit = s.getSampleImplicit(); o.sampleExplicit = s.getSampleExplicit();
    o.useEnvMap = s.getUseEnvMap(); o.useAreaLight = s.getUseAreaLight();
    o.tonemap = s.getTonemap();
    const CameraSettings c = s.getCameraSettings();
    const float pos[3] = {c.pos.x, c.pos.y, c.pos.z}, right[3] = {c.right.x, c.right.y, c.right.z}, up[3] = {c.up.x, c.up.y, c.up.z}, dir[3] = {c.dir.x, c.dir.y, c.dir.z};
    for (int k = 0; k < 3; k++) { o.camera.pos[k] = pos[k]; o.camera.right[k] = right[k]; o.camera.up[k] = up[k]; o.camera.dir[k] = dir[k]; }
    o.camera.fov = c.fov; o.camera.apertureSize = c.apertureSize; o.camera.focalDist = c.focalDist;
    const AreaLightSettings l = s.getAreaLightSettings();
    const float lr[3] = {l.right.x, l.right.y, l.right.z}, lu[3] = {l.up.x, l.up.y, l.up.z}, ln[4] = {l.N.x, l.N.y, l.N.z, l.N.w},
                lp[4] = {l.pos.x, l.pos.y, l.pos.z, l.pos.w}, le[3] = {l.E.x, l.E.y, l.E.z};
    for (int k = 0; k < 3; k++) { o.areaLight.right[k] = lr[k]; o.areaLight.up[k] = lu[k]; o.areaLight.E[k] = le[k]; }
    for (int k = 0; k < 4; k++) { o.areaLight.N[k] = ln[k]; o.areaLight.pos[k] = lp[k]; }
    o.areaLight.size[0] = l.size.x; o.areaLight.size[1] = l.size.y;
    o.parsedByReference = true;
}

json readJson(const std::string &path)
{
    std::ifstream in(path);
    if (!in) throw std::runtime_error("cannot open " + path);
    json j;
    try { in >> j; }
    catch (const std::exception &e) { throw std::runtime_error(path + ": " + e.what()); }
    return j;
}

} // namespace

namespace flx {

void loadSettingsWithReference(Settings &out, const std::string &path, const std::string &section)
{
    const json j = readJson(path);
    try
    {
        if (json_contains(j, section)) referenceSettings().import(j[section]);
#ifdef _DEBUG
        if (json_contains(j, "debug")) referenceSettings().import(j["debug"]);
#endif
    }
    catch (const std::exception &e) { throw std::runtime_error(path + ": " + e.what()); }
    snapshot(out);
}

std::vector<BenchmarkScene> loadBenchmarkFileWithReference(const std::string &path)
{
    const std::string SETTINGS_KEY = "settings", SKIP_PP_KEY = "skipPP";
    json base = readJson(path);
    const std::string baseFolder = getUnixFolderPath(path, true);
    bool skipPostProcess = false;
    auto preprocess = [&](json &j) // envMap paths are relative to the benchmark file
    {
        if (!json_contains(j, SETTINGS_KEY)) return;
        json &st = j[SETTINGS_KEY];
        if (json_contains(st, "envMap") && !isAbsolutePath(st["envMap"].get<std::string>())) st["envMap"] = baseFolder + st["envMap"].get<std::string>();
    };
    auto import = [&](const json &j)
    {
        if (!json_contains(j, SETTINGS_KEY)) return;
        const json &st = j[SETTINGS_KEY];
        referenceSettings().import(st);
        if (json_contains(st, SKIP_PP_KEY)) skipPostProcess = st[SKIP_PP_KEY].get<bool>();
    };
    std::vector<BenchmarkScene> out;
    try
    {
        referenceSettings().setRenderScale(1.0f); // src/tracer.cpp:587
        preprocess(base);
        import(base);
        std::string outputFolder = baseFolder;
        if (json_contains(base, "outputFolder"))
        {
            const std::string f = getUnixFolderPath(base["outputFolder"].get<std::string>(), false);
            outputFolder = isAbsolutePath(f) ? f : baseFolder + f;
        }
        createPath(outputFolder);
        if (!json_contains(base, "scenes")) throw std::runtime_error("no \"scenes\" array");
        for (json sceneJson : base["scenes"])
        {
            preprocess(sceneJson);
            import(base);
            import(sceneJson);
            BenchmarkScene b;
            b.file = baseFolder + sceneJson["file"].get<std::string>();
            b.outputPrefix = outputFolder + sceneJson["outputFile"].get<std::string>();
            b.skipPostProcess = skipPostProcess;
            snapshot(b.settings);
            out.push_back(b);
        }
        import(base); // "importDefaultSettings()" after the last scene (src/tracer.cpp:870)
    }
    catch (const std::runtime_error &) { throw; }
    catch (const std::exception &e) { throw std::runtime_error(path + ": " + e.what()); }
    return out;
}

} // namespace flx
