/*
 * mtscuda.h -- C ABI of libmtscuda, the B200-native back end of the `cudapath`
 * Mitsuba 0.6 integrator plugin.
 *
 * This is the drop-in boundary of the hot path: the `cudapath`
 * plugin (plugin/cudapath.cpp, an `Integrator` registered with
 * MTS_EXPORT_PLUGIN exactly like src/integrators/path/path.cpp:316-320 of the
 * reference) flattens the loaded scene and drives the renderer through the
 * entry points below.  Plain C: pointers, sizes and POD structs only; no
 * Mitsuba, torch or C++ types.  All host buffers are owned by the caller and are
 * copied during the call.  Every function returns 0 on success and a negative
 * MTS_E* code on failure; mtscuda_last_error() gives the message (the plugin
 * turns it into Log(EError, ...), src/libcore/logger.cpp:147).
 *
 * The CPU oracle (oracle/, test infrastructure only) exports the same surface
 * with the prefix `mtsoracle_` so that parity tests feed identical inputs to
 * both.  `MTS_PREFIX` selects the prefix when this header is compiled.
 *
 * Each entry cites the reference interface it replaces (paths relative to the
 * reference root).
 */
#ifndef MTSCUDA_H
#define MTSCUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef MTS_PREFIX
#define MTS_PREFIX mtscuda_
#endif
#define MTS_CAT2(a, b) a##b
#define MTS_CAT(a, b) MTS_CAT2(a, b)
#define MTS_FN(name) MTS_CAT(MTS_PREFIX, name)

#if defined(__GNUC__)
#define MTS_API __attribute__((visibility("default")))
#else
#define MTS_API
#endif

/* ---- status codes ------------------------------------------------------- */
enum {
    MTS_OK = 0,
    MTS_EINVAL = -1,      /* bad argument / unsupported feature */
    MTS_ECUDA = -2,       /* CUDA runtime failure */
    MTS_ESTATE = -3,      /* call order violated (e.g. render before build) */
    MTS_ECANCELLED = -4   /* render() observed *cancel != 0 */
};

/* ---- constants shared with the reference (include/mitsuba/core/constants.h:28-31) */
#define MTS_EPSILON 1e-4f
#define MTS_SHADOW_EPSILON 1e-3f
#define MTS_DELTA_EPSILON 1e-3f
#define MTS_FILTER_RESOLUTION 31 /* include/mitsuba/core/rfilter.h:28 */

/* ---- BSDF models (src/bsdfs/) ---------------------------------------- */
enum {
    MTS_BSDF_DIFFUSE = 0,        /* diffuse.cpp */
    MTS_BSDF_CONDUCTOR = 1,      /* conductor.cpp */
    MTS_BSDF_ROUGHCONDUCTOR = 2, /* roughconductor.cpp */
    MTS_BSDF_PLASTIC = 3,        /* plastic.cpp */
    MTS_BSDF_ROUGHPLASTIC = 4,   /* roughplastic.cpp */
    MTS_BSDF_DIELECTRIC = 5,     /* dielectric.cpp */
    MTS_BSDF_COATING = 6,        /* coating.cpp */
    MTS_BSDF_BLEND = 7,          /* blendbsdf.cpp */
    MTS_BSDF_BUMPMAP = 8,        /* bumpmap.cpp */
    MTS_BSDF_TWOSIDED = 9,       /* twosided.cpp */
    MTS_BSDF_THINDIELECTRIC = 10, /* thindielectric.cpp */
    MTS_BSDF_MASK = 11,          /* mask.cpp */
    MTS_BSDF_ROUGHDIELECTRIC = 12, /* roughdielectric.cpp */
    MTS_BSDF_MIXTURE = 13,       /* mixturebsdf.cpp */
    MTS_BSDF_ROUGHCOATING = 14,  /* roughcoating.cpp */
    MTS_BSDF_NORMALMAP = 15      /* normalmap.cpp */
};
#define MTS_MIXTURE_MAX 16 /* nested BSDFs of one mixturebsdf node */

/* node.flags */
enum {
    MTS_BF_GGX = 1,            /* microfacet.h:48-56: 0 = Beckmann, 1 = GGX */
    MTS_BF_SAMPLE_VISIBLE = 2, /* microfacet.h sampleVisible */
    MTS_BF_NONLINEAR = 4,      /* plastic.cpp / roughplastic.cpp `nonlinear` */
    MTS_BF_RT_2D = 8,          /* roughplastic: alpha is textured -> 2-D rough-transmittance slice */
    MTS_BF_RAY_DIFF = 16       /* BSDF::usesRayDifferentials() of this node as its configure() computes it (e.g.
                                  diffuse.cpp:101, coating.cpp:172-174, blendbsdf.cpp:130, bumpmap.cpp:96).  Read on the ROOT
                                  node of a shape's BSDF: Intersection::getBSDF(ray) computes the uv partials of a camera-ray
                                  hit only then (include/mitsuba/render/records.inl:69-75), and with partials present every
                                  ewa / trilinear texture of the tree is looked up filtered (src/librender/texture.cpp:104-114) */
};

/* BSDF lobe-type bits, same values as BSDF::EBSDFType (include/mitsuba/render/bsdf.h) */
enum {
    MTS_ENULL = 0x00001,
    MTS_EDIFFUSE_REFLECTION = 0x00002,
    MTS_EDIFFUSE_TRANSMISSION = 0x00004,
    MTS_EGLOSSY_REFLECTION = 0x00008,
    MTS_EGLOSSY_TRANSMISSION = 0x00010,
    MTS_EDELTA_REFLECTION = 0x00020,
    MTS_EDELTA_TRANSMISSION = 0x00040,
    MTS_EFRONTSIDE = 0x08000,
    MTS_EBACKSIDE = 0x10000
};
/* Third random number of the per-function BSDF probes (bsdf_sample takes two): BSDFs that draw from
 * bRec.sampler (roughdielectric.cpp:405) get this integer hash of the two given numbers, bit-identical in the
 * reference probe, the oracle and the CUDA library.  The renderers use their own sample streams. */
#define MTS_PROBE_U3(xbits, ybits, out) do { uint32_t h_ = (xbits) * 0x9E3779B1u ^ (ybits) * 0x85EBCA77u; h_ ^= h_ >> 15; \
    h_ *= 0xC2B2AE3Du; h_ ^= h_ >> 13; (out) = (float)(h_ >> 8) * (1.0f / 16777216.0f); } while (0)
#define MTS_EALL 0x001FF /* BSDF::EAll: every diffuse, glossy, delta and 1-D delta lobe bit */
#define MTS_ESMOOTH (MTS_EDIFFUSE_REFLECTION | MTS_EDIFFUSE_TRANSMISSION | MTS_EGLOSSY_REFLECTION | MTS_EGLOSSY_TRANSMISSION)
#define MTS_EDELTA (MTS_ENULL | MTS_EDELTA_REFLECTION | MTS_EDELTA_TRANSMISSION)
#define MTS_ETRANSMISSION (MTS_EDIFFUSE_TRANSMISSION | MTS_EGLOSSY_TRANSMISSION | MTS_EDELTA_TRANSMISSION | MTS_ENULL)

/* EMeasure (include/mitsuba/render/common.h) */
enum { MTS_MEASURE_SOLID_ANGLE = 1, MTS_MEASURE_DISCRETE = 4 };

/* A texture slot: `tex < 0` -> the constant colour `scale`
 * (ConstantSpectrumTexture / ConstantFloatTexture, include/mitsuba/hw/basicshader.h:40,97);
 * otherwise bitmap texture `tex` multiplied by `scale`
 * (ScalingTexture, src/textures/scale.cpp:52-100, folded in by the host). */
typedef struct MtsTexRef {
    float scale[3];
    int32_t tex;
} MtsTexRef;

/* One node of a flattened BSDF tree.  Children are indices into the node array.
 *
 * slot use per model:
 *   diffuse        t0 = reflectance
 *   conductor      t0 = specularReflectance; eta,k (RGB, already divided by extEta)
 *   roughconductor t0 = specularReflectance, t1 = alphaU, t2 = alphaV; eta,k
 *   plastic        t0 = diffuseReflectance, t1 = specularReflectance; eta[0] = intIOR/extIOR;
 *                  f[0] = fdrInt, f[1] = fdrExt, f[2] = specularSamplingWeight
 *   roughplastic   t0 = diffuseReflectance, t1 = specularReflectance, t2 = alpha; eta[0];
 *                  f[2] = specularSamplingWeight, f[0] = alphaMin, f[1] = alphaMax of the LUT,
 *                  f[3] = internal Fdr (1-D case); lut[0] = external table offset in the float pool,
 *                  lut[1] = thetaSamples, lut[2] = alphaSamples, lut[3] = internal diffuse table offset (2-D case)
 *   dielectric     t0 = specularReflectance, t1 = specularTransmittance; eta[0]
 *   coating        t0 = sigmaA, t1 = specularReflectance; eta[0]; f[0] = thickness,
 *                  f[2] = specularSamplingWeight; child[0] = nested
 *   blendbsdf      t0 = weight; child[0], child[1]
 *   bumpmap        t0 = displacement; child[0] = nested
 *   thindielectric t0 = specularReflectance, t1 = specularTransmittance; eta[0]
 *   mask           t0 = opacity; child[0] = nested
 *   roughdielectric t0 = specularReflectance, t1 = alphaU, t2 = alphaV; eta[0] = intIOR/extIOR;
 *                  specularTransmittance = k[0..2] (scale) x bitmap texture lut[0] (-1: constant) -- a fourth
 *                  MtsTexRef folded into the otherwise unused k/lut slots
 *   twosided       child[0] = front-side BRDF, child[1] = back-side BRDF (= child[0] when only one is given)
 *   mixturebsdf    lut[1] = N nested BSDFs (1..MTS_MIXTURE_MAX), lut[0] = offset of 3N + 1 floats in the float pool:
 *                  N node indices (stored as floats), N weights (m_weights after the energy-conservation rescale),
 *                  N + 1 entries of the normalised DiscreteDistribution CDF (m_pdf, include/mitsuba/core/pmf.h)
 *   roughcoating   t0 = sigmaA, t1 = specularReflectance, t2 = alpha; eta[0]; f[0] = alphaMin, f[1] = alphaMax of the
 *                  LUT, f[2] = specularSamplingWeight, f[3] = thickness; lut[0..2] = rough-transmittance slice as for
 *                  roughplastic (external table only); child[0] = nested
 *   normalmap      t0 = normals; child[0] = nested
 */
typedef struct MtsBsdfNode {
    int32_t type;
    int32_t flags;
    int32_t child[2];
    MtsTexRef t0, t1, t2;
    float eta[3];
    float k[3];
    float f[4];
    int32_t lut[4];
    uint32_t lobes;  /* union of the EBSDFType bits of the whole subtree (BSDF::getType()) */
    uint32_t pad0;
} MtsBsdfNode;

/* ---- textures (src/textures/bitmap.cpp, include/mitsuba/render/mipmap.h) ---- */
enum { MTS_WRAP_REPEAT = 0, MTS_WRAP_MIRROR = 1, MTS_WRAP_CLAMP = 2, MTS_WRAP_ZERO = 3, MTS_WRAP_ONE = 4 };
/* EMIPFilterType (include/mitsuba/render/mipmap.h:53-62).  TRILINEAR and EWA make the library build the MIP pyramid
 * (2-lobe Lanczos resampling of the float image, every level rounded to half; mipmap.h:238-262) and filter the
 * lookups of camera-ray hits (TMIPMap::eval, mipmap.h:629-715; evalEWA :764-834); all other lookups are bilinear at
 * level 0 (src/textures/bitmap.cpp:431-454). */
enum { MTS_FILTER_NEAREST = 0, MTS_FILTER_BILINEAR = 1, MTS_FILTER_TRILINEAR = 2, MTS_FILTER_EWA = 3 };
#define MTS_MAX_MIP_LEVELS 17 /* 65536 texels per side */

typedef struct MtsTextureDesc {
    int32_t width, height;
    int32_t channels;       /* 1 or 3; texels are float32, row-major, y down; stored as half like TMIPMap */
    int32_t wrap_u, wrap_v; /* mipmap.h:503-563 */
    int32_t filter;         /* bitmap.cpp:431-454: unfiltered lookups are bilinear (or box for `nearest`) */
    float uv_scale[2];      /* Texture2D uscale/vscale, src/librender/texture.cpp:84-94 */
    float uv_offset[2];
    float max_anisotropy;   /* bitmap.cpp:230-233: `maxAnisotropy` (default 20) for ewa, forced to 1 otherwise */
    int32_t texel_format;   /* MTS_TEXEL_F32 (0): `texels` are float32; MTS_TEXEL_F16: IEEE half, the storage type of TMIPMap
                               (mipmap.h:143-160) -- what Texture::getBitmap() of a live reference texture hands out */
} MtsTextureDesc;
enum { MTS_TEXEL_F32 = 0, MTS_TEXEL_F16 = 1 };

/* ---- sensor (src/sensors/perspective.cpp:126-178, 271-297) ------------------ */
typedef struct MtsCamera {
    float sample_to_camera[16]; /* row-major 4x4, m_sampleToCamera */
    float camera_to_world[16];  /* row-major 4x4, m_worldTransform */
    float dx[3], dy[3];         /* near-plane position differentials m_dx, m_dy */
    float near_clip, far_clip;
    float inv_resolution[2];    /* 1 / crop size, src/librender/sensor.cpp:104-108 */
} MtsCamera;

/* ---- film + reconstruction filter (include/mitsuba/render/imageblock.h:144-204,
 *      src/libcore/rfilter.cpp:37-56) */
typedef struct MtsFilm {
    int32_t width, height;                         /* crop size in pixels */
    float filter_radius;
    float filter_table[MTS_FILTER_RESOLUTION + 1]; /* normalised m_values[], last entry 0 */
    int32_t has_alpha;                             /* Film::hasAlpha(): also accumulate w * alpha, alpha = 1 where the camera ray
                                                      hit a surface, else 0 (RadianceQueryRecord::rayIntersect with EOpacity,
                                                      include/mitsuba/render/records.inl:121-137; integrator.cpp:157-160) */
} MtsFilm;

/* ---- shapes ---------------------------------------------------------------- */
typedef struct MtsMeshDesc {
    const float *positions;   /* nverts*3 world space */
    const float *normals;     /* nverts*3 or NULL (face normals) */
    const float *texcoords;   /* nverts*2 or NULL */
    const uint32_t *indices;  /* ntris*3 */
    const float *tangents;    /* ntris*6 (dpdu, dpdv; TriMesh::computeUVTangents, src/librender/trimesh.cpp:683-743) or NULL */
    uint32_t nverts, ntris;
    int32_t bsdf;             /* root node index */
    int32_t uv_tangents;      /* 1: `tangents` is ignored and TriMesh::computeUVTangents (trimesh.cpp:683-743) runs on the
                                 device from positions/texcoords -- saves the 24 B/triangle upload; 0: use `tangents` */
} MtsMeshDesc;

/* ---- emitters --------------------------------------------------------------- */
enum { MTS_EMITTER_AREA = 0, MTS_EMITTER_POINT = 1, MTS_EMITTER_ENVMAP = 2 };

typedef struct MtsEmitterDesc {
    int32_t type;
    int32_t shape;          /* area: shape id returned by add_mesh/add_rectangle/add_sphere */
    float radiance[3];      /* area: radiance; point: intensity; envmap: unused (scale below) */
    float position[3];      /* point */
    float sampling_weight;  /* Emitter::getSamplingWeight() */
    float scale;            /* envmap `scale` */
    float to_world[16];     /* envmap emitter-to-world (row-major) */
    int32_t env_width, env_height;
    const float *env_texels; /* envmap: env_width*env_height*3 float RGB, row-major */
} MtsEmitterDesc;

/* ---- render call (MonteCarloIntegrator parameters, src/librender/integrator.cpp:190-225) */
typedef struct MtsRenderParams {
    int32_t spp;             /* samples per pixel rendered by THIS call */
    int32_t sample_offset;   /* global index of this call's first sample (multi-GPU sharding) */
    int32_t sample_stride;   /* global sample index = sample_offset + i*sample_stride */
    int32_t max_depth;       /* -1 = unbounded */
    int32_t rr_depth;
    int32_t strict_normals;
    int32_t hide_emitters;
    uint32_t seed;
    int32_t clear_film;      /* 1: zero the film first (Scene::render does film->clear()) */
    int32_t wavefront_size;  /* path slots per wavefront; 0 = default */
    int32_t instrument;      /* 1: count BVH nodes visited / primitives tested (slower; never used for timing) */
    int32_t sample_count;    /* Sampler::getSampleCount() of the whole job: camera-ray differentials are scaled by
                                1/sqrt(sample_count) (src/librender/integrator.cpp:144-145, 181).  0 = spp * sample_stride */
} MtsRenderParams;

typedef struct MtsStats {
    uint64_t camera_samples;     /* paths started */
    uint64_t normal_rays;        /* "Normal rays traced", src/librender/skdtree.cpp:46 */
    uint64_t shadow_rays;        /* "Shadow rays traced", src/librender/skdtree.cpp:47 */
    uint64_t path_length_sum;    /* "Average path length" numerator, src/integrators/path/path.cpp:290-291 */
    uint64_t kernel_launches;    /* CUDA kernels launched by this render() */
    uint64_t bvh_nodes_visited;  /* closest-hit rays; filled only by instrumented runs */
    uint64_t bvh_prims_tested;
    uint64_t bvh_nodes_visited_shadow; /* shadow rays; instrumented runs only */
    uint64_t bvh_prims_tested_shadow;
    uint64_t intersect_launches, shade_launches, shadow_launches, splat_launches;
    float render_ms;             /* device time of the whole render() (CUDA events) */
    float intersect_ms, shade_ms, shadow_ms, splat_ms, raygen_ms;
    float build_ms;              /* BVH build time of the last build() */
    float upload_ms;
    /* shade stage split (CUDA events around each kernel): shade_ms = k_shade* only (surface hits), miss_ms = k_shade_miss
     * (escaped rays), sort_ms = the material counting sort; and the units they processed */
    float sort_ms, miss_ms;
    uint64_t shaded_paths;       /* surface hits shaded, summed over bounces (= normal_rays - miss_paths) */
    uint64_t miss_paths;         /* rays that escaped the scene, summed over bounces */
} MtsStats;

typedef struct MtsContext MtsContext;

/* ---- life cycle ---------------------------------------------------------- */
/* Bumped whenever a struct or signature of this header changes.  A host compiled against another version would hand
 * the library structs of the wrong size (MtsStats is written through a caller-provided pointer), so every binding
 * compares MTS_FN(abi_version)() with the MTS_ABI_VERSION it was compiled with before the first call. */
#define MTS_ABI_VERSION 6
MTS_API int32_t MTS_FN(abi_version)(void);
MTS_API const char *MTS_FN(last_error)(void);
MTS_API int MTS_FN(create)(int device, MtsContext **out);
/* A context that renders on several GPUs of the box -- the counterpart of the reference's `-p N` worker pool
 * (src/mitsuba/mitsuba.cpp:277-327, src/librender/renderproc.cpp:118-182), behind cudapath's `devices` parameter.
 * devices[0] is the primary: it receives the scene description, and build() replicates the staged scene to the other
 * devices GPU-to-GPU (every device then builds its own acceleration structure concurrently, one host thread each).
 * render() gives device d the samples d, d+n, d+2n, ... of every pixel (the union is the single-device sample set) and
 * sums the films on the primary over NVLink; read_film / get_stats / the per-function entry points act on the primary.
 * All devices need peer access to devices[0].  n = 1 is the same as create(). */
MTS_API int MTS_FN(create_multi)(const int32_t *devices, int32_t n, MtsContext **out);
MTS_API int MTS_FN(device_count)(MtsContext *ctx);
MTS_API void MTS_FN(destroy)(MtsContext *ctx);

/* ---- scene description (what cudapath::render() extracts from Scene) */
MTS_API int MTS_FN(set_camera)(MtsContext *ctx, const MtsCamera *cam);
MTS_API int MTS_FN(set_film)(MtsContext *ctx, const MtsFilm *film);
/* returns the texture id (>= 0) */
MTS_API int MTS_FN(add_texture)(MtsContext *ctx, const MtsTextureDesc *desc, const float *texels);
MTS_API int MTS_FN(set_float_pool)(MtsContext *ctx, const float *data, int64_t n);
MTS_API int MTS_FN(set_bsdfs)(MtsContext *ctx, const MtsBsdfNode *nodes, int32_t n);
/* shapes return the shape id (>= 0).  Shape::rayIntersect / fillIntersectionRecord replacements. */
MTS_API int MTS_FN(add_mesh)(MtsContext *ctx, const MtsMeshDesc *mesh);
/* src/shapes/rectangle.cpp: object space [-1,1]^2 x {0}, row-major to_world */
MTS_API int MTS_FN(add_rectangle)(MtsContext *ctx, const float to_world[16], int32_t bsdf);
/* src/shapes/sphere.cpp: object_to_world carries rotation+translation only (scale folded into radius) */
MTS_API int MTS_FN(add_sphere)(MtsContext *ctx, const float object_to_world[16], float radius, int32_t flip_normals, int32_t bsdf);
/* emitters return the emitter index (order = Scene::m_emitters order) */
MTS_API int MTS_FN(add_emitter)(MtsContext *ctx, const MtsEmitterDesc *em);
/* ShapeKDTree::build replacement + emitter pmf (src/librender/scene.cpp:375-380) + envmap CDFs */
MTS_API int MTS_FN(build)(MtsContext *ctx);

/* ---- the hot path ---------------------------------------------------------
 * SamplingIntegrator::render + renderBlock + MIPathTracer::Li + ImageBlock::put
 * (src/librender/integrator.cpp:95-188, src/integrators/path/path.cpp:119-294). */
MTS_API int MTS_FN(render)(MtsContext *ctx, const MtsRenderParams *params, volatile const int *cancel, MtsStats *stats);
/* film as width*height*4 floats (sum w*R, sum w*G, sum w*B, sum w); develop divides by the weight
 * (src/libcore/fmtconv.cpp:955-975). */
MTS_API int MTS_FN(read_film)(MtsContext *ctx, float *out_rgbw);
/* The alpha plane (width*height floats, sum w*alpha) of a film created with has_alpha = 1: the fourth channel of
 * the ESpectrumAlphaWeight block the plugin hands to Film::put (src/films/hdrfilm.cpp:391-393). */
MTS_API int MTS_FN(read_film_alpha)(MtsContext *ctx, float *out_alpha);
/* Use caller-provided device memory (width*height float4) as the film accumulator, e.g. a torch
 * tensor that is then reduced with NCCL; NULL restores the internal buffer.  CUDA library only. */
MTS_API int MTS_FN(bind_film)(MtsContext *ctx, void *device_ptr);
/* Launch everything on a caller-owned CUDA stream (a cudaStream_t passed as void*, e.g. torch's current
 * stream so that the caller's CUDA events and NCCL calls are ordered with the render); NULL restores the
 * context's own stream.  CUDA library only. */
MTS_API int MTS_FN(set_stream)(MtsContext *ctx, void *cuda_stream);
MTS_API int MTS_FN(get_stats)(MtsContext *ctx, MtsStats *stats);

/* ---- per-function entry points (parity tests bind these; the reference-side equivalents
 *      are named in the comments) -------------------------------------------------------- */

/* Sensor::sampleRayDifferential (perspective.cpp:271-297).  sample_pos: n*2 (pixel units).
 * out: n*16 floats  o(3) mint d(3) maxt rx_d(3) pad ry_d(3) pad. */
MTS_API int MTS_FN(sample_rays)(MtsContext *ctx, int64_t n, const float *sample_pos, float *out);

/* ShapeKDTree::rayIntersect(ray, its) (shadow = 0) / rayIntersect(ray) (shadow = 1),
 * src/librender/skdtree.cpp:112-142, 207-226, including the adaptive ray epsilon.
 * rays: n*8 floats  o(3) mint d(3) maxt.
 * hits: n*4 floats  t u v pad (t = inf on miss; shadow: t = 0 occluded / inf free)
 * ids:  n*2 int32   shape id, primitive index within the shape (-1,-1 on miss; may be NULL for shadow) */
MTS_API int MTS_FN(trace)(MtsContext *ctx, int64_t n, const float *rays, int32_t shadow, float *hits, int32_t *ids);

/* mtscuda_trace without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments, kernels and results as trace(); `rays`, `hits`, `ids` may be host or device pointers.
 * The working memory is one fixed array of device memory inside the library (about 2.9 MB per GPU, held until exit):
 * at most MTS_NOALLOC_MAX_RAYS rays per call and 512 material sort keys per scene, larger requests fail with MTS_EINVAL.
 * All calls on one GPU share that array: calls that overlap on the GPU race. */
#define MTS_NOALLOC_MAX_RAYS 16384
MTS_API int MTS_FN(trace_noalloc)(MtsContext *ctx, int64_t n, const float *rays, int32_t shadow, float *hits, int32_t *ids);

/* fillIntersectionRecord (include/mitsuba/render/skdtree.h:343-428).
 * out: n*24 floats  p(3) t geoN(3) u shN(3) v dpdu(3) pad dpdv(3) pad wi(3) pad */
MTS_API int MTS_FN(hit_records)(MtsContext *ctx, int64_t n, const float *rays, float *out, int32_t *ids);

/* Surface context for BSDF queries: n*16 floats  uv(2) dudx dvdx dpdu(3) dudy dpdv(3) dvdy shN(3) hasUVPartials;
 * geometric normal = shading normal; shading frame = computeShadingFrame(shN, dpdu).  The uv partials (slots 2, 3, 7, 11)
 * are honoured when slot 15 is non-zero (Intersection::hasUVPartials); all zero = an unfiltered lookup. */
/* BSDF::eval + BSDF::pdf.  wi, wo: n*3 local directions.  out_value: n*3, out_pdf: n. */
MTS_API int MTS_FN(bsdf_eval_pdf)(MtsContext *ctx, int32_t bsdf, int64_t n, const float *surf, const float *wi,
                                  const float *wo, int32_t measure, float *out_value, float *out_pdf);
/* BSDF::sample(bRec, pdf, sample).  u: n*2.  out: n*12 floats  weight(3) pdf wo(3) eta sampledType(as float) pad(3) */
MTS_API int MTS_FN(bsdf_sample)(MtsContext *ctx, int32_t bsdf, int64_t n, const float *surf, const float *wi,
                                const float *u, float *out);

/* Scene::sampleEmitterDirect without the visibility test (src/librender/scene.cpp:828-852).
 * ref: n*3, ref_n: n*3 (0 = no orientation test), u: n*2.
 * out: n*16 floats  value(3) pdf d(3) dist p(3) measure n(3) emitter_index */
MTS_API int MTS_FN(sample_emitter_direct)(MtsContext *ctx, int64_t n, const float *ref, const float *ref_n,
                                          const float *u, float *out);
/* Emitter::evalEnvironment (envmap.cpp:380-410, no differentials) and the solid-angle
 * pdf of Scene::pdfEmitterDirect for the environment emitter.  d: n*3.  out: n*4 value(3) pdf */
MTS_API int MTS_FN(eval_environment)(MtsContext *ctx, int64_t n, const float *d, float *out);

/* Texture2D::eval(its, filter = true) of bitmap texture `tex` (src/librender/texture.cpp:104-114 ->
 * BitmapTexture::eval, src/textures/bitmap.cpp:431-454, 486-499 -> TMIPMap::eval / evalBilinear / evalBox).
 * uv: n*2 surface uv (the texture's uscale/offset are applied like Texture2D does); duv: n*4 = dudx dvdx dudy dvdy, or
 * NULL for lookups without partials.  out: n*3 */
MTS_API int MTS_FN(texture_eval)(MtsContext *ctx, int32_t tex, int64_t n, const float *uv, const float *duv, float *out);

/* Camera-ray hits with their uv partials: Sensor::sampleRayDifferential(sample_pos) + scaleDifferential(1/sqrt(
 * sample_count)) + Scene::rayIntersect + Intersection::computePartials (src/librender/intersection.cpp:5-76).
 * sample_pos: n*2 (pixel units).  out: n*8 floats  dudx dvdx dudy dvdy t hit(1/0) u v */
MTS_API int MTS_FN(hit_partials)(MtsContext *ctx, int64_t n, const float *sample_pos, int32_t sample_count, float *out);

/* Emitter::evalEnvironment(ray) for camera rays WITH differentials (envmap.cpp:380-410: EWA lookup, maxAnisotropy 10).
 * sample_pos: n*2 (pixel units).  out: n*3 */
MTS_API int MTS_FN(eval_environment_camera)(MtsContext *ctx, int64_t n, const float *sample_pos, int32_t sample_count, float *out);

/* ImageBlock::put(pos, spec, alpha) for n samples into the (cleared) film. pos: n*2, value: n*3 */
MTS_API int MTS_FN(splat)(MtsContext *ctx, int64_t n, const float *pos, const float *value, int32_t clear_first);

#ifdef __cplusplus
}
#endif
#endif /* MTSCUDA_H */
