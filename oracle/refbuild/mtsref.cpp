/*
 * mtsref.cpp -- TEST INFRASTRUCTURE ONLY: a plain-C front end over the REFERENCE's own classes (compiled unchanged
 * from /root/reference by oracle/refbuild/Makefile).  It replaces the two Xerces-based pieces of the reference,
 * src/mitsuba/mitsuba.cpp (process set-up, render job) and src/librender/scenehandler.cpp (XML -> objects), with
 * the same sequence of reference calls driven from Python (oracle/refbuild/mtsref.py):
 *
 *     Properties -> PluginManager::createObject -> addChild/setParent -> configure     (scenehandler.cpp:700-780)
 *     Scheduler + LocalWorkers, RenderQueue, RenderJob::start, waitLeft(0)             (mitsuba.cpp:277-404)
 *
 * and adds per-function probes (BSDF eval/pdf/sample, Scene::rayIntersect, sampleEmitterDirect, ...) that call the
 * reference objects on caller-supplied inputs, laid out exactly like the probes of include/mtscuda.h, so that tests
 * can put reference, oracle restatement and CUDA results side by side.  No arithmetic of the path lives here.
 */
#include <mitsuba/core/plugin.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/core/sched.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/fstream.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/appender.h>
#include <mitsuba/core/properties.h>
#include <mitsuba/render/scene.h>
#include <mitsuba/render/renderjob.h>
#include <mitsuba/render/renderqueue.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/bsdf.h>
#include <mitsuba/render/emitter.h>
#include <mitsuba/render/texture.h>
#include <mitsuba/render/sensor.h>
#include <mitsuba/render/film.h>
#include <mitsuba/render/sampler.h>
#include <mitsuba/render/integrator.h>
#include <mitsuba/render/shape.h>
#include <chrono>
#include <cstring>
#include <iostream>
#include <map>
#include <string>

using namespace mitsuba;

namespace {
thread_local std::string g_err;
bool g_init = false;
int g_threads = 0;
int fail(const std::string &m) { g_err = m; return -1; }

#define GUARD_BEGIN try {
#define GUARD_END } catch (const std::exception &e) { return fail(e.what()); } catch (...) { return fail("unknown exception"); }

const Class *classOf(const std::string &kind) {
    if (kind == "bsdf") return MTS_CLASS(BSDF);
    if (kind == "shape") return MTS_CLASS(Shape);
    if (kind == "emitter") return MTS_CLASS(Emitter);
    if (kind == "sensor") return MTS_CLASS(Sensor);
    if (kind == "film") return MTS_CLASS(Film);
    if (kind == "sampler") return MTS_CLASS(Sampler);
    if (kind == "rfilter") return MTS_CLASS(ReconstructionFilter);
    if (kind == "texture") return MTS_CLASS(Texture);
    if (kind == "integrator") return MTS_CLASS(Integrator);
    return NULL;
}

Intersection makeIts(const float *s) {
    /* same convention as the `surf` arrays of include/mtscuda.h: uv(2) dudx dvdx dpdu(3) dudy dpdv(3) dvdy n(3) hasUVPartials */
    Intersection its;
    its.p = Point(0.0f);
    its.t = 1.0f;
    its.uv = Point2(s[0], s[1]);
    its.dpdu = Vector(s[4], s[5], s[6]);
    its.dpdv = Vector(s[8], s[9], s[10]);
    Normal n(s[12], s[13], s[14]);
    its.geoFrame = Frame(n);
    its.shFrame.n = n;
    computeShadingFrame(its.shFrame.n, its.dpdu, its.shFrame);
    its.hasUVPartials = s[15] != 0.0f;
    its.dudx = s[2]; its.dvdx = s[3]; its.dudy = s[7]; its.dvdy = s[11];
    its.time = 0;
    its.shape = NULL;
    its.primIndex = 0;
    its.color = Spectrum(0.0f);
    its.wi = Vector(0, 0, 1);
    return its;
}
void rgb(const Spectrum &s, float *out) { Float r, g, b; s.toLinearRGB(r, g, b); out[0] = r; out[1] = g; out[2] = b; }
} // namespace

extern "C" {

const char *mtsref_last_error() { return g_err.c_str(); }

/* mitsuba.cpp:420-433 + :280-286.  `root` holds plugins/ and data/ (oracle/_ref). */
int mtsref_init(const char *root, int nthreads, int logLevel) {
    GUARD_BEGIN
    if (g_init) return 0;
    Class::staticInitialization();
    Object::staticInitialization();
    PluginManager::staticInitialization();
    Statistics::staticInitialization();
    Thread::staticInitialization();
    Logger::staticInitialization();
    FileStream::staticInitialization();
    Spectrum::staticInitialization();
    Bitmap::staticInitialization();
    Scheduler::staticInitialization();
    /* stdout belongs to the caller (bench.py prints one JSON line there): log to stderr, no progress bar */
    Logger *logger = Thread::getThread()->getLogger();
    logger->clearAppenders();
    logger->addAppender(new StreamAppender(&std::cerr));
    logger->setLogLevel((ELogLevel)logLevel);
    ProgressReporter::setEnabled(false);
    ref<FileResolver> fr = Thread::getThread()->getFileResolver();
    fr->appendPath(fs::path(root));
    if (nthreads <= 0) nthreads = getCoreCount();
    g_threads = nthreads;
    Scheduler *sched = Scheduler::getInstance();
    for (int i = 0; i < nthreads; ++i)
        sched->registerWorker(new LocalWorker(i, formatString("wrk%i", i)));
    sched->start();
    g_init = true;
    return 0;
    GUARD_END
}
int mtsref_thread_count() { return g_threads; }
/* mitsuba.cpp:452-463 (without SceneHandler / SHVector, which are not part of the oracle build) */
void mtsref_shutdown() {
    if (!g_init) return;
    g_init = false;
    try {
        Scheduler::getInstance()->stop();
        Scheduler::staticShutdown();
        Bitmap::staticShutdown();
        Spectrum::staticShutdown();
        FileStream::staticShutdown();
        Logger::staticShutdown();
        Thread::staticShutdown();
        Statistics::staticShutdown();
        PluginManager::staticShutdown();
        Object::staticShutdown();
        Class::staticShutdown();
    } catch (...) {}
}

/* ---- Properties ---------------------------------------------------------------------------- */
void *mtsref_props_new(const char *pluginName) { return new Properties(pluginName); }
void mtsref_props_free(void *p) { delete (Properties *)p; }
void mtsref_props_set_bool(void *p, const char *n, int v) { ((Properties *)p)->setBoolean(n, v != 0); }
void mtsref_props_set_int(void *p, const char *n, int v) { ((Properties *)p)->setInteger(n, v); }
void mtsref_props_set_float(void *p, const char *n, float v) { ((Properties *)p)->setFloat(n, v); }
void mtsref_props_set_string(void *p, const char *n, const char *v) { ((Properties *)p)->setString(n, v); }
void mtsref_props_set_point(void *p, const char *n, float x, float y, float z) { ((Properties *)p)->setPoint(n, Point(x, y, z)); }
void mtsref_props_set_vector(void *p, const char *n, float x, float y, float z) { ((Properties *)p)->setVector(n, Vector(x, y, z)); }
void mtsref_props_set_rgb(void *p, const char *n, float r, float g, float b) {
    Spectrum s; s.fromLinearRGB(r, g, b);   /* <rgb> / <spectrum> with 3 samples (scenehandler.cpp) */
    ((Properties *)p)->setSpectrum(n, s);
}
void mtsref_props_set_transform(void *p, const char *n, const float *m /* row-major 4x4 */) {
    Matrix4x4 mtx;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) mtx.m[i][j] = m[4 * i + j];
    ((Properties *)p)->setTransform(n, Transform(mtx));
}

/* ---- objects (scenehandler.cpp:700-780) ---------------------------------------------------- */
void *mtsref_create(const char *kind, void *props) {
    try {
        ConfigurableObject *o = NULL;
        Properties *pr = (Properties *)props;
        if (std::string(kind) == "scene") o = new Scene(*pr);
        else {
            const Class *c = classOf(kind);
            if (!c) { fail(std::string("unknown object kind ") + kind); return NULL; }
            o = PluginManager::getInstance()->createObject(c, *pr);
        }
        o->incRef();
        return o;
    } catch (const std::exception &e) { fail(e.what()); return NULL; }
}
int mtsref_add_child(void *parent, const char *name, void *child) {
    GUARD_BEGIN
    ((ConfigurableObject *)parent)->addChild(name, (ConfigurableObject *)child);
    ((ConfigurableObject *)child)->setParent((ConfigurableObject *)parent);
    return 0;
    GUARD_END
}
int mtsref_configure(void *o) {
    GUARD_BEGIN
    ((ConfigurableObject *)o)->configure();
    return 0;
    GUARD_END
}
void mtsref_release(void *o) { if (o) ((ConfigurableObject *)o)->decRef(); }

/* ---- scene ------------------------------------------------------------------------------------- */
int mtsref_scene_initialize(void *s, double *kdBuildSeconds) {
    GUARD_BEGIN
    auto t0 = std::chrono::steady_clock::now();
    ((Scene *)s)->initialize();   /* builds the kd-tree, the emitter pmf, ... (scene.cpp:300-400) */
    if (kdBuildSeconds) *kdBuildSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return 0;
    GUARD_END
}

/* mitsuba.cpp:375-395: RenderJob on the scheduler's workers; the film comes back as (sum w*R, sum w*G, sum w*B, sum w). */
int mtsref_render(void *s, float *outRGBW, double *seconds, const char *destFile, float *outAlpha /* may be NULL */) {
    GUARD_BEGIN
    Scene *scene = (Scene *)s;
    scene->setDestinationFile(fs::path(destFile ? destFile : "/tmp/mtsref_out"));
    Statistics::getInstance()->resetAll();
    ref<RenderQueue> queue = new RenderQueue();
    auto t0 = std::chrono::steady_clock::now();
    ref<RenderJob> job = new RenderJob("ren0", scene, queue, -1, -1, -1, true, false);
    job->start();
    queue->waitLeft(0);
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    queue = NULL;
    if (outRGBW) {
        Film *film = scene->getFilm();
        Vector2i size = film->getCropSize();
        ref<Bitmap> bmp = new Bitmap(Bitmap::ESpectrumAlphaWeight, Bitmap::EFloat32, size);
        film->develop(Point2i(0, 0), size, Point2i(0, 0), bmp);
        const float *src = bmp->getFloat32Data();
        const int ch = bmp->getChannelCount();   /* SPECTRUM_SAMPLES + 2 */
        for (size_t i = 0; i < (size_t)size.x * size.y; ++i) {
            outRGBW[4 * i] = src[ch * i]; outRGBW[4 * i + 1] = src[ch * i + 1]; outRGBW[4 * i + 2] = src[ch * i + 2];
            outRGBW[4 * i + 3] = src[ch * i + ch - 1];
            if (outAlpha) outAlpha[i] = src[ch * i + ch - 2];   /* ESpectrumAlphaWeight: ..., alpha, weight */
        }
    }
    return 0;
    GUARD_END
}
/* human-readable counters ("Normal rays traced", "Shadow rays traced", "Average path length", ...) */
int mtsref_stats(char *buf, int n) {
    GUARD_BEGIN
    std::string s = Statistics::getInstance()->getStats();
    strncpy(buf, s.c_str(), n - 1); buf[n - 1] = 0;
    return (int)s.size();
    GUARD_END
}

/* ---- probes (same array layouts as include/mtscuda.h) ---------------------------------------- */
int mtsref_bsdf_eval_pdf(void *b, int64_t n, const float *surf, const float *wi, const float *wo, int measure, float *outV, float *outP) {
    GUARD_BEGIN
    const BSDF *bsdf = (const BSDF *)b;
    for (int64_t i = 0; i < n; ++i) {
        Intersection its = makeIts(surf + 16 * i);
        BSDFSamplingRecord bRec(its, Vector(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), Vector(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), ERadiance);
        rgb(bsdf->eval(bRec, (EMeasure)measure), outV + 3 * i);
        outP[i] = bsdf->pdf(bRec, (EMeasure)measure);
    }
    return 0;
    GUARD_END
}
/* BSDFs with EUsesSampler (roughdielectric) draw one more number through bRec.sampler->next1D().  The probes feed
   it a value derived from the two given numbers by an integer hash (exact on every side; include/mtscuda.h,
   MTS_PROBE_U3), so the reference, the oracle and the CUDA library see identical inputs. */
class ProbeSampler : public Sampler {
public:
    ProbeSampler() : Sampler(Properties()), value(0) { }
    Float next1D() { return value; }
    Point2 next2D() { return Point2(value, value); }
    Float value;
};
static float probeU3(float a, float b) {
    uint32_t x, y; memcpy(&x, &a, 4); memcpy(&y, &b, 4);
    uint32_t h = x * 0x9E3779B1u ^ y * 0x85EBCA77u; h ^= h >> 15; h *= 0xC2B2AE3Du; h ^= h >> 13;
    return (float)(h >> 8) * (1.0f / 16777216.0f);
}
int mtsref_bsdf_sample(void *b, int64_t n, const float *surf, const float *wi, const float *u, float *out) {
    GUARD_BEGIN
    const BSDF *bsdf = (const BSDF *)b;
    ref<ProbeSampler> ps = new ProbeSampler();
    for (int64_t i = 0; i < n; ++i) {
        Intersection its = makeIts(surf + 16 * i);
        ps->value = probeU3(u[2 * i], u[2 * i + 1]);
        BSDFSamplingRecord bRec(its, ps.get(), ERadiance);
        bRec.wi = Vector(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]);
        Float pdf = 0;
        Spectrum w = bsdf->sample(bRec, pdf, Point2(u[2 * i], u[2 * i + 1]));
        float *q = out + 12 * i;
        for (int k = 0; k < 12; ++k) q[k] = 0;
        if (w.isZero()) continue;
        rgb(w, q); q[3] = pdf; q[4] = bRec.wo.x; q[5] = bRec.wo.y; q[6] = bRec.wo.z; q[7] = bRec.eta; q[8] = (float)bRec.sampledType;
    }
    return 0;
    GUARD_END
}
int mtsref_bsdf_type(void *b) { return (int)((const BSDF *)b)->getType(); }

static int shapeIndex(const Scene *scene, const Shape *s) {
    const ref_vector<Shape> &shapes = const_cast<Scene *>(scene)->getShapes();
    for (size_t i = 0; i < shapes.size(); ++i) if (shapes[i].get() == s) return (int)i;
    return -1;
}
int mtsref_trace(void *s, int64_t n, const float *rays, int shadow, float *hits, int32_t *ids) {
    GUARD_BEGIN
    const Scene *scene = (const Scene *)s;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7], 0.0f);
        float *q = hits + 4 * i;
        q[1] = q[2] = q[3] = 0;
        if (ids) ids[2 * i] = ids[2 * i + 1] = -1;
        if (shadow) { q[0] = scene->rayIntersect(ray) ? 0.0f : std::numeric_limits<float>::infinity(); continue; }
        Intersection its;
        if (scene->rayIntersect(ray, its)) {
            q[0] = its.t;
            if (ids) { ids[2 * i] = shapeIndex(scene, its.shape); ids[2 * i + 1] = its.shape->getClass()->derivesFrom(MTS_CLASS(TriMesh)) ? (int)its.primIndex : 0; }
        } else q[0] = std::numeric_limits<float>::infinity();
    }
    return 0;
    GUARD_END
}
int mtsref_hit_records(void *s, int64_t n, const float *rays, float *out, int32_t *ids) {
    GUARD_BEGIN
    const Scene *scene = (const Scene *)s;
    for (int64_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        Ray ray(Point(r[0], r[1], r[2]), Vector(r[4], r[5], r[6]), r[3], r[7], 0.0f);
        float *q = out + 24 * i;
        for (int k = 0; k < 24; ++k) q[k] = 0;
        if (ids) ids[2 * i] = ids[2 * i + 1] = -1;
        Intersection its;
        if (!scene->rayIntersect(ray, its)) { q[3] = std::numeric_limits<float>::infinity(); continue; }
        q[0] = its.p.x; q[1] = its.p.y; q[2] = its.p.z; q[3] = its.t;
        q[4] = its.geoFrame.n.x; q[5] = its.geoFrame.n.y; q[6] = its.geoFrame.n.z; q[7] = its.uv.x;
        q[8] = its.shFrame.n.x; q[9] = its.shFrame.n.y; q[10] = its.shFrame.n.z; q[11] = its.uv.y;
        q[12] = its.dpdu.x; q[13] = its.dpdu.y; q[14] = its.dpdu.z;
        q[16] = its.dpdv.x; q[17] = its.dpdv.y; q[18] = its.dpdv.z;
        q[20] = its.wi.x; q[21] = its.wi.y; q[22] = its.wi.z;
        if (ids) { ids[2 * i] = shapeIndex(scene, its.shape); ids[2 * i + 1] = its.shape->getClass()->derivesFrom(MTS_CLASS(TriMesh)) ? (int)its.primIndex : 0; }
    }
    return 0;
    GUARD_END
}
/* Scene::sampleEmitterDirect(dRec, sample, testVisibility = false) (scene.cpp:828-852) */
int mtsref_sample_emitter_direct(void *s, int64_t n, const float *ref_, const float *refn, const float *u, float *out) {
    GUARD_BEGIN
    const Scene *scene = (const Scene *)s;
    const ref_vector<Emitter> &ems = const_cast<Scene *>(scene)->getEmitters();
    for (int64_t i = 0; i < n; ++i) {
        DirectSamplingRecord dRec(Point(ref_[3 * i], ref_[3 * i + 1], ref_[3 * i + 2]), 0.0f);
        dRec.refN = Normal(refn[3 * i], refn[3 * i + 1], refn[3 * i + 2]);
        Spectrum v = scene->sampleEmitterDirect(dRec, Point2(u[2 * i], u[2 * i + 1]), false);
        float *q = out + 16 * i;
        for (int k = 0; k < 16; ++k) q[k] = 0;
        q[15] = -1;
        if (v.isZero() && dRec.pdf == 0) continue;
        rgb(v, q); q[3] = dRec.pdf; q[4] = dRec.d.x; q[5] = dRec.d.y; q[6] = dRec.d.z; q[7] = dRec.dist;
        q[8] = dRec.p.x; q[9] = dRec.p.y; q[10] = dRec.p.z; q[11] = (float)dRec.measure; q[12] = dRec.n.x; q[13] = dRec.n.y; q[14] = dRec.n.z;
        int idx = -1;
        for (size_t k = 0; k < ems.size(); ++k) if (ems[k].get() == dRec.object) idx = (int)k;
        q[15] = (float)idx;
    }
    return 0;
    GUARD_END
}
/* Emitter::evalEnvironment without differentials + Scene::pdfEmitterDirect for the environment emitter */
int mtsref_eval_environment(void *s, int64_t n, const float *d, float *out) {
    GUARD_BEGIN
    const Scene *scene = (const Scene *)s;
    const Emitter *env = scene->getEnvironmentEmitter();
    if (!env) return fail("no environment emitter");
    for (int64_t i = 0; i < n; ++i) {
        Vector dir(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        RayDifferential ray(Point(0.0f), dir, 0.0f);
        ray.hasDifferentials = false;
        rgb(env->evalEnvironment(ray), out + 4 * i);
        DirectSamplingRecord dRec(Point(0.0f), 0.0f);
        dRec.refN = Normal(0.0f);
        if (env->fillDirectSamplingRecord(dRec, ray)) { dRec.object = env; out[4 * i + 3] = scene->pdfEmitterDirect(dRec); }
        else out[4 * i + 3] = 0;
    }
    return 0;
    GUARD_END
}
/* Texture::eval(its, filter = true) of a texture object, with or without uv partials */
int mtsref_texture_eval(void *t, int64_t n, const float *uv, const float *duv, float *out) {
    GUARD_BEGIN
    const Texture *tex = (const Texture *)t;
    const float zero[16] = {0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    for (int64_t i = 0; i < n; ++i) {
        Intersection its = makeIts(zero);
        its.uv = Point2(uv[2 * i], uv[2 * i + 1]);
        if (duv) { its.hasUVPartials = true; its.dudx = duv[4 * i]; its.dvdx = duv[4 * i + 1]; its.dudy = duv[4 * i + 2]; its.dvdy = duv[4 * i + 3]; }
        rgb(tex->eval(its, true), out + 3 * i);
    }
    return 0;
    GUARD_END
}
/* What SamplingIntegrator::renderBlock hands to Li (integrator.cpp:144-145, 178-181) */
static void cameraRay(const Sensor *sensor, const float *sp, int sampleCount, RayDifferential &ray) {
    sensor->sampleRayDifferential(ray, Point2(sp[0], sp[1]), Point2(0.5f), 0.0f);
    ray.scaleDifferential(1.0f / std::sqrt((Float)sampleCount));
}
/* Scene::rayIntersect + Intersection::computePartials for camera rays.  out: n*8 = dudx dvdx dudy dvdy t hit u v */
int mtsref_hit_partials(void *s, int64_t n, const float *sp, int sampleCount, float *out) {
    GUARD_BEGIN
    const Scene *scene = (const Scene *)s;
    for (int64_t i = 0; i < n; ++i) {
        RayDifferential ray; cameraRay(scene->getSensor(), sp + 2 * i, sampleCount, ray);
        Intersection its;
        float *q = out + 8 * i;
        for (int k = 0; k < 8; ++k) q[k] = 0;
        if (!scene->rayIntersect(ray, its)) continue;
        its.computePartials(ray);
        q[0] = its.dudx; q[1] = its.dvdx; q[2] = its.dudy; q[3] = its.dvdy; q[4] = its.t; q[5] = 1.0f; q[6] = its.uv.x; q[7] = its.uv.y;
    }
    return 0;
    GUARD_END
}
/* Emitter::evalEnvironment of camera rays WITH differentials (envmap.cpp:380-410) */
int mtsref_eval_environment_camera(void *s, int64_t n, const float *sp, int sampleCount, float *out) {
    GUARD_BEGIN
    const Scene *scene = (const Scene *)s;
    const Emitter *env = scene->getEnvironmentEmitter();
    if (!env) return fail("no environment emitter");
    for (int64_t i = 0; i < n; ++i) {
        RayDifferential ray; cameraRay(scene->getSensor(), sp + 2 * i, sampleCount, ray);
        rgb(env->evalEnvironment(ray), out + 3 * i);
    }
    return 0;
    GUARD_END
}
/* Sensor::sampleRayDifferential (perspective.cpp:271-297) */
int mtsref_sample_rays(void *s, int64_t n, const float *sp, float *out) {
    GUARD_BEGIN
    const Sensor *sensor = ((Scene *)s)->getSensor();
    for (int64_t i = 0; i < n; ++i) {
        RayDifferential ray;
        sensor->sampleRayDifferential(ray, Point2(sp[2 * i], sp[2 * i + 1]), Point2(0.5f), 0.0f);
        float *q = out + 16 * i;
        q[0] = ray.o.x; q[1] = ray.o.y; q[2] = ray.o.z; q[3] = ray.mint; q[4] = ray.d.x; q[5] = ray.d.y; q[6] = ray.d.z; q[7] = ray.maxt;
        q[8] = ray.rxDirection.x; q[9] = ray.rxDirection.y; q[10] = ray.rxDirection.z; q[11] = 0;
        q[12] = ray.ryDirection.x; q[13] = ray.ryDirection.y; q[14] = ray.ryDirection.z; q[15] = 0;
    }
    return 0;
    GUARD_END
}
/* ImageBlock::put(pos, spec, alpha) with the film's reconstruction filter into a full-frame block */
int mtsref_splat(void *s, int64_t n, const float *pos, const float *value, float *outRGBW) {
    GUARD_BEGIN
    Film *film = ((Scene *)s)->getFilm();
    Vector2i size = film->getCropSize();
    ref<ImageBlock> block = new ImageBlock(Bitmap::ESpectrumAlphaWeight, size, film->getReconstructionFilter());
    block->setOffset(Point2i(0, 0));
    block->clear();
    for (int64_t i = 0; i < n; ++i) {
        Spectrum v; v.fromLinearRGB(value[3 * i], value[3 * i + 1], value[3 * i + 2]);
        block->put(Point2(pos[2 * i], pos[2 * i + 1]), v, 1.0f);
    }
    const Bitmap *bmp = block->getBitmap();
    const int border = block->getBorderSize(), ch = bmp->getChannelCount(), bw = bmp->getWidth();
    const float *src = bmp->getFloat32Data();
    for (int y = 0; y < size.y; ++y) for (int x = 0; x < size.x; ++x) {
        const float *p = src + ((size_t)(y + border) * bw + (x + border)) * ch;
        float *q = outRGBW + 4 * ((size_t)y * size.x + x);
        q[0] = p[0]; q[1] = p[1]; q[2] = p[2]; q[3] = p[ch - 1];
    }
    return 0;
    GUARD_END
}

} // extern "C"
