// flx::Settings — the values of the reference's `Settings` singleton (reference src/settings.hpp:29-113) as a plain
// struct the headless host passes around, with the defaults of Settings::init (src/settings.cpp:17-59).
//
// Reading settings.json / a .bm.json:
//   * built with the reference tree (FLX_WITH_REFERENCE_HOST): the reference's OWN src/settings.cpp is compiled in place
//     and does the parsing — Settings::import with all its keys (scalars, "camera" with pos / dir / lookAt / fov /
//     apertureSize / focalDist / cameraRotation, "areaLight" with pos / N / E / size) — and this struct is a snapshot of
//     the singleton's getters (ref_settings.cpp);
//   * built without it: a small scanner for the scalar keys of the "release" object (camera / areaLight keep defaults).
#pragma once
#include <cctype>
#include <cstdint>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace flx {

struct Settings
{
    std::string envMapName;
    float renderScale = 1.0f;
    int windowWidth = 640, windowHeight = 480;
    uint32_t wfBufferSize = 1u << 20;
    bool useWavefront = false; // the reference's default renderer is the microkernel one; fluctus_headless picks with --renderer
    bool useRussianRoulette = false, useSeparateQueues = false, clUseBitstack = false;
    int maxPathDepth = 10;
    uint32_t maxSpp = 0, maxRenderTime = 0;
    bool sampleImplicit = true, sampleExplicit = true, useEnvMap = false, useAreaLight = true;
    int tonemap = 2;
    // CameraSettings / AreaLightSettings (src/settings.hpp:8-27), defaults src/settings.cpp:40-58
    struct Camera { float pos[3] = {0.0f, 1.0f, 3.5f}, right[3] = {1.0f, 0.0f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f}, dir[3] = {0.0f, 0.0f, -1.0f};
                    float fov = 60.0f, apertureSize = 0.0f, focalDist = 0.5f; } camera;
    struct AreaLight { float right[3] = {0.0f, 0.0f, -1.0f}, up[3] = {0.0f, 1.0f, 0.0f}, N[4] = {-1.0f, 0.0f, 0.0f, 0.0f}, pos[4] = {1.0f, 1.0f, 0.0f, 1.0f};
                       float E[3] = {100.0f, 100.0f, 100.0f}, size[2] = {0.5f, 0.5f}; } areaLight;
    bool parsedByReference = false; // which of the two readers filled this struct

    // settings.json: the "release" object (Settings::load, src/settings.cpp:61-87)
    void load(const std::string &path, const std::string &section = "release");

private:
    void scan(const std::string &path, const std::string &section);
    static bool find(const std::string &obj, const std::string &key, std::string &out)
    {
        const size_t k = obj.find("\"" + key + "\"");
        if (k == std::string::npos) return false;
        size_t p = obj.find(':', k);
        if (p == std::string::npos) return false;
        p++;
        while (p < obj.size() && isspace((unsigned char)obj[p])) p++;
        if (p >= obj.size()) return false;
        if (obj[p] == '"') { const size_t e = obj.find('"', p + 1); out = obj.substr(p + 1, e - p - 1); return true; }
        size_t e = p;
        while (e < obj.size() && obj[e] != ',' && obj[e] != '}' && !isspace((unsigned char)obj[e])) e++;
        out = obj.substr(p, e - p);
        return !out.empty();
    }
};

// One entry of a benchmark file (Tracer::runBenchmarkFromFile, src/tracer.cpp:567-700): the scene to load, the prefix
// of its output files (<outputFolder>/<outputFile> + .ppm / .csv / .txt) and the settings in force for it — the
// file's "settings" object imported first, the scene's own "settings" on top, as the reference does per scene.
struct BenchmarkScene
{
    std::string file, outputPrefix;
    Settings settings;
    bool skipPostProcess = false;
};

#ifdef FLX_WITH_REFERENCE_HOST
// ref_settings.cpp — on top of the reference's own Settings::import
void loadSettingsWithReference(Settings &out, const std::string &path, const std::string &section);
std::vector<BenchmarkScene> loadBenchmarkFileWithReference(const std::string &path);
#endif

inline void Settings::load(const std::string &path, const std::string &section)
{
#ifdef FLX_WITH_REFERENCE_HOST
    loadSettingsWithReference(*this, path, section);
#else
    scan(path, section);
#endif
}

// A deliberately small scanner: `"key": value` pairs with number / bool / string values, nested objects ignored.
inline void Settings::scan(const std::string &path, const std::string &section)
{
    std::ifstream in(path);
    if (!in) throw std::runtime_error("cannot open settings file " + path);
    std::stringstream ss; ss << in.rdbuf();
    const std::string all = ss.str();
    size_t a = all.find("\"" + section + "\"");
    if (a == std::string::npos) return;
    a = all.find('{', a);
    int depth = 0; size_t b = a;
    for (; b < all.size(); b++) { if (all[b] == '{') depth++; else if (all[b] == '}' && --depth == 0) break; }
    const std::string obj = all.substr(a, b - a);
    std::string s;
    if (find(obj, "envMap", s)) envMapName = s;
    if (find(obj, "renderScale", s)) renderScale = std::stof(s);
    if (find(obj, "windowWidth", s)) windowWidth = std::stoi(s);
    if (find(obj, "windowHeight", s)) windowHeight = std::stoi(s);
    if (find(obj, "wfBufferSize", s)) wfBufferSize = (uint32_t)std::stoul(s);
    if (find(obj, "useWavefront", s)) useWavefront = s == "true";
    if (find(obj, "useRussianRoulette", s)) useRussianRoulette = s == "true";
    if (find(obj, "useSeparateQueues", s)) useSeparateQueues = s == "true";
    if (find(obj, "clUseBitstack", s)) clUseBitstack = s == "true";
    if (find(obj, "maxPathDepth", s)) maxPathDepth = std::stoi(s);
    if (find(obj, "maxSpp", s)) maxSpp = (uint32_t)std::stoul(s);
    if (find(obj, "maxRenderTime", s)) maxRenderTime = (uint32_t)std::stoul(s);
    if (find(obj, "sampleImplicit", s)) sampleImplicit = s == "true";
    if (find(obj, "sampleExplicit", s)) sampleExplicit = s == "true";
    if (find(obj, "useEnvMap", s)) useEnvMap = s == "true";
    if (find(obj, "useAreaLight", s)) useAreaLight = s == "true";
    if (find(obj, "tonemap", s)) tonemap = std::stoi(s);
}

} // namespace flx
