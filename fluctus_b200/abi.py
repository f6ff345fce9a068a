"""numpy views of the boundary structs in include/fluctus_abi.h, and the RenderParams builder that
mirrors how the reference's Tracer fills them (src/tracer.cpp:43-65 resetParams, :1127-1168
initCamera / initPostProcessing / initAreaLight, defaults of src/settings.cpp:17-59)."""
import math

import numpy as np

TRI_DT = np.dtype([("v", np.float32, (3, 3, 4)), ("matId", np.int32), ("_pad", np.int32, 3)])  # [vertex][p,n,t][xyzw]
NODE_DT = np.dtype([("bmin", np.float32, 4), ("bmax", np.float32, 4), ("parent", np.int32),
                    ("iStartOrRight", np.uint32), ("nPrims", np.uint8), ("_pad", np.uint8, 7)])
MAT_DT = np.dtype([("Kd", np.float32, 4), ("Ks", np.float32, 4), ("Ke", np.float32, 4), ("Kt", np.float32, 4),
                   ("Ns", np.float32), ("Ni", np.float32), ("d", np.float32), ("map_Kd", np.int32),
                   ("map_Ks", np.int32), ("map_N", np.int32), ("type", np.int32), ("_pad", np.int32)])
TEX_DT = np.dtype([("offset", np.uint32), ("width", np.uint32), ("height", np.uint32)])
HIT_DT = np.dtype([("P", np.float32, 4), ("N", np.float32, 4), ("uv", np.float32, 2), ("t", np.float32),
                   ("i", np.int32), ("areaLightHit", np.int32), ("matId", np.int32), ("_pad", np.int32, 2)])
COUNTERS_DT = np.dtype([(n, np.uint32) for n in ("raygenQueue", "extensionQueue", "shadowQueue", "diffuseQueue",
                                                  "glossyQueue", "ggxReflQueue", "ggxRefrQueue", "deltaQueue",
                                                  "emissiveQueue", "splattedSamples")])
AREA_LIGHT_DT = np.dtype([("right", np.float32, 4), ("up", np.float32, 4), ("N", np.float32, 4), ("pos", np.float32, 4),
                          ("E", np.float32, 4), ("size", np.float32, 2), ("_pad", np.float32, 2)])
CAMERA_DT = np.dtype([("pos", np.float32, 4), ("dir", np.float32, 4), ("up", np.float32, 4), ("right", np.float32, 4),
                      ("fov", np.float32), ("fovSCALE", np.float32), ("apertureSize", np.float32), ("focalDist", np.float32)])
PARAMS_DT = np.dtype([("areaLight", AREA_LIGHT_DT), ("camera", CAMERA_DT), ("exposure", np.float32), ("tmOperator", np.uint32),
                      ("width", np.uint32), ("height", np.uint32), ("n_tris", np.uint32), ("useEnvMap", np.uint32),
                      ("useAreaLight", np.uint32), ("envMapStrength", np.float32), ("maxBounces", np.uint32),
                      ("sampleImpl", np.uint32), ("sampleExpl", np.uint32), ("useRoulette", np.uint32),
                      ("wfSeparateQueues", np.uint32), ("maxSpp", np.uint32), ("worldRadius", np.float32),
                      ("width1", np.float32), ("height1", np.float32), ("_pad", np.uint32, 3)])
assert TRI_DT.itemsize == 160 and NODE_DT.itemsize == 48 and MAT_DT.itemsize == 96 and TEX_DT.itemsize == 12
assert HIT_DT.itemsize == 64 and COUNTERS_DT.itemsize == 40 and AREA_LIGHT_DT.itemsize == 96 and CAMERA_DT.itemsize == 80
assert PARAMS_DT.itemsize == 256 and PARAMS_DT.fields["width"][1] == 184 and PARAMS_DT.fields["worldRadius"][1] == 232

# reference src/bxdf_types.h:4-12
BXDF_DIFFUSE, BXDF_GLOSSY, BXDF_GGX_ROUGH_REFLECTION, BXDF_IDEAL_REFLECTION = 1 << 1, 1 << 2, 1 << 3, 1 << 4
BXDF_GGX_ROUGH_DIELECTRIC, BXDF_IDEAL_DIELECTRIC, BXDF_EMISSIVE, BXDF_MIXED = 1 << 5, 1 << 6, 1 << 7, 1 << 8

# slot numbers of the reference's SoA GPUTaskState (include/fluctus_abi.h)
SLOTS = dict(orig=0, dir=4, shadowOrig=8, shadowDir=12, T=16, Ei=20, lastBsdf=24, lastEmission=28, lastT=32, P=36, N=40,
             uvTex=44, phase=46, lastPdfW=47, pathLen=48, seed=49, lastSpecular=50, shadowRayBlocked=51, backfaceHit=52,
             pixelIndex=53, firstDiffuseHit=54, lastPdfDirect=55, lastPdfImplicit=56, lastCosTh=57,
             lastLightPickProb=58, shadowRayLen=59, t=60, i=61, areaLightHit=62, matId=63)
SLOT_WIDTH = dict(orig=3, dir=3, shadowOrig=3, shadowDir=3, T=3, Ei=3, lastBsdf=3, lastEmission=3, lastT=3, P=3, N=3, uvTex=2)
INT_SLOTS = {"pathLen", "seed", "lastSpecular", "shadowRayBlocked", "backfaceHit", "pixelIndex", "firstDiffuseHit",
             "i", "areaLightHit", "matId", "phase"}
QUEUES = ("raygen", "extension", "shadow", "diffuse", "glossy", "ggxRefl", "ggxRefr", "delta", "emissive")


def _v4(x):
    return np.array([x[0], x[1], x[2], 0.0], np.float32)


def make_render_params(width, height, *, n_tris=0, world_radius=1.0, max_bounces=10, max_spp=0,
                       use_env_map=False, use_area_light=True, env_map_strength=1.0,
                       sample_implicit=True, sample_explicit=True, use_roulette=False, separate_queues=False,
                       tonemap=2, exposure=1.0,
                       cam_pos=(0.0, 1.0, 3.5), cam_dir=(0.0, 0.0, -1.0), cam_up=(0.0, 1.0, 0.0), cam_right=(1.0, 0.0, 0.0),
                       fov=60.0, aperture=0.0, focal_dist=0.5,
                       light_pos=(1.0, 1.0, 0.0), light_n=(-1.0, 0.0, 0.0), light_right=(0.0, 0.0, -1.0),
                       light_up=(0.0, 1.0, 0.0), light_e=(100.0, 100.0, 100.0), light_size=(0.5, 0.5)):
    """RenderParams with the reference's defaults (settings.cpp:17-59); returns a 1-element PARAMS_DT array."""
    p = np.zeros(1, PARAMS_DT)
    r = p[0]
    r["areaLight"]["right"] = _v4(light_right)
    r["areaLight"]["up"] = _v4(light_up)
    r["areaLight"]["N"] = _v4(light_n)
    r["areaLight"]["pos"] = _v4(light_pos)
    r["areaLight"]["E"] = _v4(light_e)
    r["areaLight"]["size"] = np.array(light_size, np.float32)
    r["camera"]["pos"] = _v4(cam_pos)
    r["camera"]["dir"] = _v4(cam_dir)
    r["camera"]["up"] = _v4(cam_up)
    r["camera"]["right"] = _v4(cam_right)
    r["camera"]["fov"] = fov
    # cam.fovSCALE = tan(toRad(0.5f * cam.fov)) in float (tracer.cpp:1137)
    r["camera"]["fovSCALE"] = np.float32(math.tan(float(np.float32(np.float32(0.5) * np.float32(fov)) * np.float32(0.01745329251994329576923690768489))))
    r["camera"]["apertureSize"] = aperture
    r["camera"]["focalDist"] = focal_dist
    r["exposure"] = exposure
    r["tmOperator"] = tonemap
    r["width"], r["height"], r["n_tris"] = width, height, n_tris
    r["useEnvMap"], r["useAreaLight"] = int(use_env_map), int(use_area_light)
    r["envMapStrength"] = env_map_strength
    r["maxBounces"] = max_bounces
    r["sampleImpl"], r["sampleExpl"] = int(sample_implicit), int(sample_explicit)
    r["useRoulette"], r["wfSeparateQueues"] = int(use_roulette), int(separate_queues)
    r["maxSpp"] = max_spp
    r["worldRadius"] = world_radius
    r["width1"] = np.float32(1.0) / np.float32(width)
    r["height1"] = np.float32(1.0) / np.float32(height)
    return p


def world_radius_of(bounds):
    """params.worldRadius = length(bounds.max - bounds.min) * 0.5f (tracer.cpp:79-80), in float."""
    d = (np.asarray(bounds[1], np.float32) - np.asarray(bounds[0], np.float32)).astype(np.float32)
    s = np.float32(d[0] * d[0]) + np.float32(d[1] * d[1])
    s = np.float32(s) + np.float32(d[2] * d[2])
    return float(np.float32(np.sqrt(np.float32(s))) * np.float32(0.5))


def state_field(state64, name):
    """Named field of an exported (64, N) uint32 state array (reference SoA layout)."""
    s = SLOTS[name]
    w = SLOT_WIDTH.get(name, 1)
    raw = state64[s:s + w]
    out = raw if name in INT_SLOTS else raw.view(np.float32)
    if name in ("i", "areaLightHit", "matId"):
        out = raw.view(np.int32)
    return out[0] if w == 1 else out.T


def state_record(params):
    """The viewer's state file data/states/state_<hash>.dat as Tracer::iterateStateItems writes it (src/tracer.cpp:996-1054):
    44 little-endian words — camera rotation / speed (viewer-only, zero here), fov, focalDist, apertureSize, dir, pos,
    right, up, area light N / pos / right / up / E / size, envMapStrength, maxBounces, useAreaLight, useEnvMap, sampleExpl,
    sampleImpl, useRoulette, exposure, tonemap operator.  `fluctus_headless --load-state` reads it."""
    r = params[0] if getattr(params, "ndim", 0) else params
    f = np.zeros(44, np.float32)
    u = f.view(np.uint32)
    cam, al = r["camera"], r["areaLight"]
    f[3], f[4], f[5] = cam["fov"], cam["focalDist"], cam["apertureSize"]
    f[6:9], f[9:12], f[12:15], f[15:18] = cam["dir"][:3], cam["pos"][:3], cam["right"][:3], cam["up"][:3]
    f[18:21], f[21:24], f[24:27], f[27:30], f[30:33] = al["N"][:3], al["pos"][:3], al["right"][:3], al["up"][:3], al["E"][:3]
    f[33], f[34], f[35] = al["size"][0], al["size"][1], r["envMapStrength"]
    u[36], u[37], u[38], u[39], u[40], u[41] = (r["maxBounces"], r["useAreaLight"], r["useEnvMap"], r["sampleExpl"],
                                                r["sampleImpl"], r["useRoulette"])
    f[42] = r["exposure"]
    u[43] = r["tmOperator"]
    return f
