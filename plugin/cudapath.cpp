/*
 * cudapath -- drop-in Mitsuba 0.6 integrator plugin that renders with the B200-native path tracer (libmtscuda).
 *
 * Build it inside a Mitsuba 0.6 tree as src/integrators/cudapath/cudapath.cpp (SConscript line:
 *     plugins += env.SharedLibrary('cudapath', ['cudapath/cudapath.cpp'], LIBS = env['LIBS'] + ['mtscuda'])
 * ) and use it from an unchanged scene file with <integrator type="cudapath"> -- same parameters as `path`
 * (maxDepth, rrDepth, strictNormals, hideEmitters; src/librender/integrator.cpp:190-225).  In this repository it is
 * compiled against the reference sources by oracle/refbuild/Makefile and exercised through the reference's own
 * PluginManager (tests/test_gpu_plugin.py).
 *
 * The class is an `Integrator` (not a `SamplingIntegrator`: there is no per-ray Li()).  render() flattens the
 * configured object graph into the plain-C scene description of include/mtscuda.h, renders, and hands the result to
 * the Film as one ESpectrumAlphaWeight ImageBlock (Film::put), i.e. through the same develop path as the CPU
 * integrators.  Plugin classes such as SmoothDiffuse or BitmapTexture are private to their .so files and have no
 * getters, so the material graph is read through the mechanism Mitsuba itself uses to ship scenes to render nodes:
 * it is serialized into a MemoryStream (every class writes its configured members in a fixed order, e.g.
 * src/bsdfs/roughplastic.cpp serialize()) and that stream is parsed here.  Scenes using plugins outside the scope of
 * the path are rejected with Log(EError, ...), never rendered differently.
 */
#include <mitsuba/mitsuba.h>
/* InstanceManager keeps the object <-> id tables of a serialization private; the flattener needs them to get from an id in
   the stream back to the live Texture (whose texels it then reads through Texture::getBitmap() instead of decoding the
   image file a second time).  Access only -- the class is compiled inside libmitsuba-core as it is. */
#if defined(__MITSUBA_CORE_SERIALIZATION_H_)
#error "serialization.h must not have been included yet"
#endif
#define private public
#include <mitsuba/core/serialization.h>
#undef private
#include <mitsuba/render/scene.h>
#include <mitsuba/render/renderqueue.h>
#include <mitsuba/render/renderjob.h>
#include <mitsuba/render/imageblock.h>
#include <mitsuba/render/trimesh.h>
#include <mitsuba/core/mstream.h>
#include <mitsuba/core/bitmap.h>
#include <mitsuba/core/statistics.h>
#include <mitsuba/core/fresolver.h>
#include <mitsuba/core/fstream.h>
#include <map>
#include <vector>
#include <cstdlib>
#include <sys/time.h>

#include "mtscuda.h"

/* RoughTransmittance (the precomputed transmittance tables of roughplastic) is a header of the bsdfs directory */
#define protected public
#include "microfacet.h"
#include "rtrans.h"
#undef protected

MTS_NAMESPACE_BEGIN

static StatsCounter statLaunches("cudapath", "CUDA kernel launches");
/* the same counters the CPU path feeds (src/librender/skdtree.cpp:46-47, src/integrators/path/path.cpp:24), so that
   `mitsuba -v` / Statistics::getStats() report the GPU render like a CPU one */
static StatsCounter statNormalRays("General", "Normal rays traced");
static StatsCounter statShadowRays("General", "Shadow rays traced");
static StatsCounter statAvgPathLength("Path tracer", "Average path length", EAverage);

/* The stream the material graph is serialized into.  BitmapTexture::serialize (src/textures/bitmap.cpp) appends the whole
   image FILE to the stream, through Stream::copyTo's 512-byte blocks (src/libcore/stream.cpp:428-441) -- a gigabyte of
   copying for the textures of one scene that the flattener never looks at (it reads the texels from the live texture).
   This stream keeps everything except those blocks: a write of exactly 512 bytes is counted, not stored, and the parser
   skips the correspondingly shorter blob (dropped()). */
class FieldStream : public MemoryStream {
public:
    FieldStream() : MemoryStream(4096), m_inRun(false), m_lastStored(0) {}
    void write(const void *ptr, size_t size) {
        /* a blob follows its 8-byte length (writeSize); anything else of 512 bytes is an ordinary field */
        if (size == 512 && (m_inRun || m_lastStored == 8)) { m_dropped[getPos()] += size; m_inRun = true; return; }
        m_inRun = false; m_lastStored = size;
        MemoryStream::write(ptr, size);
    }
    /* bytes dropped at stored position `pos` */
    size_t dropped(size_t pos) const { std::map<size_t, size_t>::const_iterator it = m_dropped.find(pos); return it == m_dropped.end() ? 0 : it->second; }
private:
    std::map<size_t, size_t> m_dropped;
    bool m_inRun;
    size_t m_lastStored;
};

class CudaPath : public Integrator {
public:
    CudaPath(const Properties &props) : Integrator(props), m_cancel(0) {
        /* identical to MonteCarloIntegrator (src/librender/integrator.cpp:190-225) */
        m_rrDepth = props.getInteger("rrDepth", 5);
        m_maxDepth = props.getInteger("maxDepth", -1);
        m_strictNormals = props.getBoolean("strictNormals", false);
        m_hideEmitters = props.getBoolean("hideEmitters", false);
        m_device = props.getInteger("device", 0);          /* optional extras must be queried */
        /* `devices`: how many GPUs of the box share the render, starting at `device` (the analogue of `mitsuba -p N`):
           each takes every N-th sample of every pixel, the films are summed over NVLink (mtscuda_create_multi) */
        m_devices = props.getInteger("devices", 1);
        if (m_devices < 1 || m_devices > 16)
            Log(EError, "'devices' must be between 1 and 16!");
        m_seed = (uint32_t) props.getInteger("seed", 0);
        /* `exactTextures` (default true; false, or MTS_CUDAPATH_FAST_TEXTURES=1 in the environment, selects the fast route):
           build the MIP pyramids from the decoded image FILES like the reference does, instead of from level 0 of the
           reference's own pyramid (see parseTexture) */
        m_exactTextures = props.getBoolean("exactTextures", getenv("MTS_CUDAPATH_FAST_TEXTURES") == NULL);
        if (m_rrDepth <= 0)
            Log(EError, "'rrDepth' must be set to a value greater than zero!");
        if (m_maxDepth <= 0 && m_maxDepth != -1)
            Log(EError, "'maxDepth' must be set to -1 (infinite) or a value greater than zero!");
    }

    CudaPath(Stream *stream, InstanceManager *manager) : Integrator(stream, manager), m_cancel(0) {
        m_rrDepth = stream->readInt();
        m_maxDepth = stream->readInt();
        m_strictNormals = stream->readBool();
        m_hideEmitters = stream->readBool();
        m_device = stream->readInt();
        m_devices = stream->readInt();
        m_seed = stream->readUInt();
        m_exactTextures = stream->readBool();
    }

    void serialize(Stream *stream, InstanceManager *manager) const {
        Integrator::serialize(stream, manager);
        stream->writeInt(m_rrDepth);
        stream->writeInt(m_maxDepth);
        stream->writeBool(m_strictNormals);
        stream->writeBool(m_hideEmitters);
        stream->writeInt(m_device);
        stream->writeInt(m_devices);
        stream->writeUInt(m_seed);
        stream->writeBool(m_exactTextures);
    }

    void cancel() { m_cancel = 1; }   /* may arrive from another thread (scene.cpp:451-456) */

    bool render(Scene *scene, RenderQueue *queue, const RenderJob *job, int sceneResID, int sensorResID, int samplerResID) {
        m_cancel = 0;
        MtsContext *ctx = NULL;
        /* MTS_PLUGIN_TIMING=1: wall-clock of the stages of this call on stderr */
        const bool timing = getenv("MTS_PLUGIN_TIMING") != NULL;
        struct Clock { static double now() { timeval tv; gettimeofday(&tv, NULL); return tv.tv_sec + 1e-6 * tv.tv_usec; } };
        const double tStart = Clock::now();
        double tCreate = 0, tFlatten = 0, tBuild = 0, tRender = 0;
        if (mtscuda_abi_version() != MTS_ABI_VERSION)
            Log(EError, "libmtscuda implements ABI version %i, this plugin was compiled against version %i -- rebuild both "
                "from the same include/mtscuda.h", (int) mtscuda_abi_version(), (int) MTS_ABI_VERSION);
        if (m_devices > 1) {
            std::vector<int32_t> ids(m_devices);
            for (int i = 0; i < m_devices; ++i) ids[i] = m_device + i;
            check(NULL, mtscuda_create_multi(&ids[0], m_devices, &ctx));
        } else
            check(NULL, mtscuda_create(m_device, &ctx));
        bool ok = false;
        try {
            tCreate = Clock::now();
            Flattener f(this, ctx);
            f.run(scene);
            tFlatten = Clock::now();
            check(ctx, mtscuda_build(ctx));
            tBuild = Clock::now();

            Film *film = scene->getFilm();
            const Vector2i size = film->getCropSize();
            MtsRenderParams p;
            memset(&p, 0, sizeof(p));
            p.spp = (int32_t) scene->getSampler()->getSampleCount();
            p.sample_offset = 0; p.sample_stride = 1;
            p.max_depth = m_maxDepth; p.rr_depth = m_rrDepth;
            p.strict_normals = m_strictNormals; p.hide_emitters = m_hideEmitters;
            p.seed = m_seed; p.clear_film = 1;
            p.sample_count = p.spp;                      /* scales the camera-ray differentials (integrator.cpp:144-145) */
            Log(EInfo, "Starting render job (%ix%i, " SIZE_T_FMT " samples, %i CUDA device%s from %i) ..", size.x, size.y,
                scene->getSampler()->getSampleCount(), m_devices, m_devices > 1 ? "s" : "", m_device);
            MtsStats st;
            int rc = mtscuda_render(ctx, &p, &m_cancel, &st);
            tRender = Clock::now();
            if (rc == MTS_ECANCELLED) {
                Log(EWarn, "Rendering was cancelled");
            } else {
                check(ctx, rc);
                statLaunches += st.kernel_launches;
                statNormalRays += st.normal_rays;
                statShadowRays += st.shadow_rays;
                statAvgPathLength.incrementBase(st.camera_samples);
                statAvgPathLength += st.path_length_sum;
                Log(EInfo, "cudapath: %.1f ms on the device, %.1f Msamples/s, %.1f Mrays/s (BVH build %.1f ms)",
                    st.render_ms, st.camera_samples / (st.render_ms * 1e3), (st.normal_rays + st.shadow_rays) / (st.render_ms * 1e3), st.build_ms);
                /* film: (sum w*R, sum w*G, sum w*B, sum w) -> one ESpectrumAlphaWeight block (hdrfilm.cpp:391-393) */
                std::vector<float> rgbw((size_t) size.x * size.y * 4), alpha;
                check(ctx, mtscuda_read_film(ctx, rgbw.data()));
                if (film->hasAlpha()) {   /* sum w * alpha, alpha = 0 where the camera ray escaped (records.inl:121-137) */
                    alpha.resize((size_t) size.x * size.y);
                    check(ctx, mtscuda_read_film_alpha(ctx, alpha.data()));
                }
                ref<ImageBlock> block = new ImageBlock(Bitmap::ESpectrumAlphaWeight, size, film->getReconstructionFilter());
                /* block offsets are relative to the crop window (renderproc.cpp bindResource starts at (0, 0)); the
                   crop offset itself is already part of the camera's sample-to-camera transform */
                block->setOffset(Point2i(0, 0));
                block->clear();
                Bitmap *bmp = block->getBitmap();
                const int border = block->getBorderSize(), ch = bmp->getChannelCount(), bw = bmp->getWidth();
                Float *dst = bmp->getFloatData();
                for (int y = 0; y < size.y; ++y) for (int x = 0; x < size.x; ++x) {
                    const float *s = &rgbw[4 * ((size_t) y * size.x + x)];
                    Float *d = dst + ((size_t) (y + border) * bw + (x + border)) * ch;
                    Spectrum spec; spec.fromLinearRGB(s[0], s[1], s[2]);
                    for (int k = 0; k < SPECTRUM_SAMPLES; ++k) d[k] = spec[k];
                    d[SPECTRUM_SAMPLES] = alpha.empty() ? s[3] : alpha[(size_t) y * size.x + x];   /* without EOpacity alpha stays 1: w * 1 */
                    d[SPECTRUM_SAMPLES + 1] = s[3];   /* weight */
                }
                film->put(block);
                ok = true;
                if (timing)
                    fprintf(stderr, "[cudapath] flatten: serialize %.0f ms, getBitmap %.0f ms, texel statistics %.0f ms, add_texture %.0f ms, add_mesh %.0f ms\n",
                            f.m_tSerialize * 1e3, f.m_tGetBitmap * 1e3, f.m_tConvert * 1e3, f.m_tAddTexture * 1e3, f.m_tMesh * 1e3);
                if (timing)
                    fprintf(stderr, "[cudapath] create %.0f ms, flatten + upload %.0f ms, build %.0f ms, render %.0f ms, film %.0f ms\n",
                            (tCreate - tStart) * 1e3, (tFlatten - tCreate) * 1e3, (tBuild - tFlatten) * 1e3, (tRender - tBuild) * 1e3, (Clock::now() - tRender) * 1e3);
            }
        } catch (...) {
            mtscuda_destroy(ctx);
            throw;
        }
        mtscuda_destroy(ctx);
        return ok;
    }

    std::string toString() const {
        std::ostringstream oss;
        oss << "CudaPath[" << endl << "  maxDepth = " << m_maxDepth << "," << endl << "  rrDepth = " << m_rrDepth << ","
            << endl << "  strictNormals = " << m_strictNormals << "," << endl << "  device = " << m_device << "," << endl << "  devices = " << m_devices << endl << "]";
        return oss.str();
    }

    MTS_DECLARE_CLASS()

private:
    /* libmtscuda error -> Log(EError) (throws std::runtime_error, src/libcore/logger.cpp:147) */
    void check(MtsContext *, int rc) const {
        if (rc < 0)
            Log(EError, "libmtscuda: %s", mtscuda_last_error());
    }

    /* ------------------------------------------------------------------------------------------------------- */
    struct TexInfo { MtsTexRef ref; Spectrum min, max, avg; bool constant; };

    class Flattener {
    public:
        Flattener(const CudaPath *owner, MtsContext *ctx) : m_owner(owner), m_ctx(ctx), m_tSerialize(0), m_tGetBitmap(0), m_tConvert(0), m_tAddTexture(0), m_tMesh(0) {}
        static double now() { timeval tv; gettimeofday(&tv, NULL); return tv.tv_sec + 1e-6 * tv.tv_usec; }

        void run(const Scene *scene) {
            sensorAndFilm(scene);
            /* shapes in Scene::m_shapes order; area emitters are appended after the scene-level emitters, which is
               the order of Scene::m_emitters (scene.cpp:322-336, 497-516) */
            const ref_vector<Shape> &shapes = const_cast<Scene *>(scene)->getShapes();
            std::vector<std::pair<const Emitter *, int> > areaEmitters;
            std::vector<int> shapeIds(shapes.size());
            std::vector<int> shapeBsdf(shapes.size());
            for (size_t i = 0; i < shapes.size(); ++i) {
                const Shape *s = shapes[i].get();
                if (s->hasSubsurface() || s->isMediumTransition())
                    err("subsurface scattering and participating media are out of scope for cudapath");
                shapeBsdf[i] = bsdf(s->getBSDF());
            }
            /* all textures / BSDF nodes are known now */
            if (!m_pool.empty()) m_owner->check(m_ctx, mtscuda_set_float_pool(m_ctx, m_pool.data(), (int64_t) m_pool.size()));
            m_owner->check(m_ctx, mtscuda_set_bsdfs(m_ctx, m_nodes.data(), (int32_t) m_nodes.size()));
            for (size_t i = 0; i < shapes.size(); ++i) {
                shapeIds[i] = shape(shapes[i].get(), shapeBsdf[i]);
                if (shapes[i]->isEmitter()) areaEmitters.push_back(std::make_pair(shapes[i]->getEmitter(), shapeIds[i]));
            }
            const ref_vector<Emitter> &emitters = const_cast<Scene *>(scene)->getEmitters();
            for (size_t i = 0; i < emitters.size(); ++i) {
                const Emitter *e = emitters[i].get();
                int sid = -1;
                for (size_t k = 0; k < areaEmitters.size(); ++k) if (areaEmitters[k].first == e) sid = areaEmitters[k].second;
                emitter(e, sid);
            }
        }

    private:
        void err(const std::string &msg) const { SLog(EError, "cudapath: %s", msg.c_str()); }

        static void toRowMajor(const Transform &t, float *out) {
            const Matrix4x4 &m = t.getMatrix();
            for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) out[4 * i + j] = (float) m.m[i][j];
        }

        /* ---- sensor + film (perspective.cpp:126-178, film.cpp, rfilter.h:76-77) ---- */
        void sensorAndFilm(const Scene *scene) {
            const Sensor *sensor = scene->getSensor();
            if (sensor->getClass()->getName() != "PerspectiveCameraImpl" && !sensor->getClass()->derivesFrom(MTS_CLASS(PerspectiveCamera)))
                err("a `perspective` sensor is required (sensor class " + sensor->getClass()->getName() + " is out of scope)");
            const PerspectiveCamera *cam = static_cast<const PerspectiveCamera *>(sensor);
            const Film *film = sensor->getFilm();
            const Vector2i filmSize = film->getSize(), cropSize = film->getCropSize();
            const Point2i cropOffset = film->getCropOffset();
            const Float aspect = sensor->getAspect();
            Vector2 relSize((Float) cropSize.x / (Float) filmSize.x, (Float) cropSize.y / (Float) filmSize.y);
            Point2 relOffset((Float) cropOffset.x / (Float) filmSize.x, (Float) cropOffset.y / (Float) filmSize.y);
            Transform cameraToSample =
                  Transform::scale(Vector(1.0f / relSize.x, 1.0f / relSize.y, 1.0f))
                * Transform::translate(Vector(-relOffset.x, -relOffset.y, 0.0f))
                * Transform::scale(Vector(-0.5f, -0.5f * aspect, 1.0f))
                * Transform::translate(Vector(-1.0f, -1.0f / aspect, 0.0f))
                * Transform::perspective(cam->getXFov(), cam->getNearClip(), cam->getFarClip());
            Transform sampleToCamera = cameraToSample.inverse();
            Vector2 invRes(1.0f / cropSize.x, 1.0f / cropSize.y);
            Vector dx = sampleToCamera(Point(invRes.x, 0.0f, 0.0f)) - sampleToCamera(Point(0.0f));
            Vector dy = sampleToCamera(Point(0.0f, invRes.y, 0.0f)) - sampleToCamera(Point(0.0f));
            MtsCamera c;
            toRowMajor(sampleToCamera, c.sample_to_camera);
            toRowMajor(cam->getWorldTransform()->eval(0), c.camera_to_world);
            for (int i = 0; i < 3; ++i) { c.dx[i] = dx[i]; c.dy[i] = dy[i]; }
            c.near_clip = cam->getNearClip(); c.far_clip = cam->getFarClip();
            c.inv_resolution[0] = invRes.x; c.inv_resolution[1] = invRes.y;
            m_owner->check(m_ctx, mtscuda_set_camera(m_ctx, &c));

            const ReconstructionFilter *rf = film->getReconstructionFilter();
            MtsFilm mf;
            mf.width = cropSize.x; mf.height = cropSize.y; mf.filter_radius = rf->getRadius();
            for (int i = 0; i <= MTS_FILTER_RESOLUTION; ++i)   /* m_values[i] through the public accessor */
                mf.filter_table[i] = rf->evalDiscretized((i + 0.5f) * rf->getRadius() / MTS_FILTER_RESOLUTION);
            mf.filter_table[MTS_FILTER_RESOLUTION] = 0;
            mf.has_alpha = film->hasAlpha() ? 1 : 0;
            m_owner->check(m_ctx, mtscuda_set_film(m_ctx, &mf));
            if (scene->getSampler()->getClass()->getName() != "IndependentSampler")
                err("sampler class " + scene->getSampler()->getClass()->getName() + " is out of scope (independent)");
        }

        /* ---- shapes ---- */
        int shape(const Shape *s, int bsdfIndex) {
            const std::string cls = s->getClass()->getName();
            const Properties &props = s->getProperties();
            if (cls == "Rectangle") {
                Transform tw = props.getTransform("toWorld", Transform());
                if (props.getBoolean("flipNormals", false)) tw = tw * Transform::scale(Vector(1, 1, -1));
                float m[16]; toRowMajor(tw, m);
                int id = mtscuda_add_rectangle(m_ctx, m, bsdfIndex);
                m_owner->check(m_ctx, id);
                return id;
            } else if (cls == "Sphere") {
                /* sphere.cpp:108-121: objectToWorld = toWorld * scale(1/r) * translate(center), radius *= r */
                Transform o2w = Transform::translate(Vector(props.getPoint("center", Point(0.0f))));
                Float radius = props.getFloat("radius", 1.0f);
                if (props.hasProperty("toWorld")) {
                    Transform toWorld = props.getTransform("toWorld");
                    Float r = toWorld(Vector(1, 0, 0)).length();
                    o2w = toWorld * Transform::scale(Vector(1 / r)) * o2w;
                    radius *= r;
                }
                float m[16]; toRowMajor(o2w, m);
                int id = mtscuda_add_sphere(m_ctx, m, radius, props.getBoolean("flipNormals", false) ? 1 : 0, bsdfIndex);
                m_owner->check(m_ctx, id);
                return id;
            } else if (s->getClass()->derivesFrom(MTS_CLASS(TriMesh))) {
                const TriMesh *mesh = static_cast<const TriMesh *>(s);
                MtsMeshDesc d;
                memset(&d, 0, sizeof(d));
                d.positions = (const float *) mesh->getVertexPositions();
                d.normals = mesh->hasVertexNormals() ? (const float *) mesh->getVertexNormals() : NULL;
                d.texcoords = mesh->hasVertexTexcoords() ? (const float *) mesh->getVertexTexcoords() : NULL;
                d.indices = (const uint32_t *) mesh->getTriangles();
                /* the reference already holds TriMesh::computeUVTangents' output; the library recomputes the same
                   values on the device (bit-identical, no FMA contraction) instead of uploading 24 B/triangle */
                d.tangents = NULL;
                d.uv_tangents = (mesh->hasUVTangents() && d.texcoords) ? 1 : 0;
                d.nverts = (uint32_t) mesh->getVertexCount(); d.ntris = (uint32_t) mesh->getTriangleCount();
                d.bsdf = bsdfIndex;
                const double tM = now();
                int id = mtscuda_add_mesh(m_ctx, &d);
                m_tMesh += now() - tM;
                m_owner->check(m_ctx, id);
                return id;
            }
            err("shape class " + cls + " is out of scope (rectangle, sphere, triangle meshes)");
            return -1;
        }

        /* ---- emitters ---- */
        void emitter(const Emitter *e, int shapeId) {
            const std::string cls = e->getClass()->getName();
            const Properties &props = e->getProperties();
            MtsEmitterDesc d;
            memset(&d, 0, sizeof(d));
            d.shape = -1; d.sampling_weight = e->getSamplingWeight(); d.scale = 1;
            toRowMajor(Transform(), d.to_world);
            ref<Bitmap> bmp;
            if (cls == "AreaLight") {
                if (shapeId < 0) err("area emitter without a shape");
                d.type = MTS_EMITTER_AREA; d.shape = shapeId;
                Float r, g, b; props.getSpectrum("radiance", Spectrum::getD65()).toLinearRGB(r, g, b);
                d.radiance[0] = r; d.radiance[1] = g; d.radiance[2] = b;
            } else if (cls == "PointEmitter") {
                d.type = MTS_EMITTER_POINT;
                Float r, g, b; (props.getSpectrum("intensity", Spectrum::getD65())).toLinearRGB(r, g, b);
                d.radiance[0] = r; d.radiance[1] = g; d.radiance[2] = b;
                Point p = e->getWorldTransform()->eval(0)(Point(0.0f));
                d.position[0] = p.x; d.position[1] = p.y; d.position[2] = p.z;
            } else if (cls == "EnvironmentMap") {
                d.type = MTS_EMITTER_ENVMAP;
                d.scale = props.getFloat("scale", 1.0f);
                toRowMajor(e->getWorldTransform()->eval(0), d.to_world);
                bmp = e->getBitmap()->convert(Bitmap::ERGB, Bitmap::EFloat32, 1.0f);   /* level 0 of the MIP map */
                d.env_width = bmp->getWidth(); d.env_height = bmp->getHeight();
                d.env_texels = bmp->getFloat32Data();
            } else
                err("emitter class " + cls + " is out of scope (area, point, envmap)");
            m_owner->check(m_ctx, mtscuda_add_emitter(m_ctx, &d));
        }

        /* ---- materials: parse the serialized object graph ---- */
        int bsdf(const BSDF *b) {
            if (!b) err("shape without a BSDF");
            std::map<const BSDF *, int>::iterator it = m_bsdfIds.find(b);
            if (it != m_bsdfIds.end()) return it->second;
            ref<FieldStream> ms = new FieldStream();
            ref<InstanceManager> mgr = new InstanceManager();
            const double tS = now();
            mgr->serialize(ms, b);
            m_tSerialize += now() - tS;
            ms->seek(0);
            m_idToObj.clear();                  /* id in the stream -> live object of this serialization */
            for (std::map<const SerializableObject *, unsigned int>::const_iterator i2 = mgr->m_objToId.begin(); i2 != mgr->m_objToId.end(); ++i2)
                m_idToObj[i2->second] = i2->first;
            std::map<uint32_t, int> seen;
            std::map<uint32_t, TexInfo> tex;   /* object ids are unique per InstanceManager, i.e. per root BSDF */
            int idx = parseBsdf(ms, seen, tex);
            /* only the root's flag is read by the library (Intersection::getBSDF(ray), records.inl:69-75) */
            if (b->usesRayDifferentials()) m_nodes[idx].flags |= MTS_BF_RAY_DIFF;
            m_bsdfIds[b] = idx;
            return idx;
        }

        static Spectrum lum3(Float v) { return Spectrum(v); }

        TexInfo constant(const Spectrum &v) {
            TexInfo t; Float r, g, b; v.toLinearRGB(r, g, b);
            t.ref.scale[0] = r; t.ref.scale[1] = g; t.ref.scale[2] = b; t.ref.tex = -1;
            t.min = t.max = t.avg = v; t.constant = true;
            return t;
        }

        TexInfo parseTexture(Stream *s, std::map<uint32_t, TexInfo> &seen) {
            uint32_t id = s->readUInt();
            if (id == 0) err("missing texture");
            if (seen.count(id)) return seen[id];
            std::string cls = s->readString();
            TexInfo t;
            if (cls == "ConstantSpectrumTexture") t = constant(Spectrum(s));
            else if (cls == "ConstantFloatTexture") t = constant(Spectrum(s->readFloat()));
            else if (cls == "ScalingTexture") {          /* src/textures/scale.cpp:147-151 */
                t = parseTexture(s, seen);
                Spectrum sc(s);
                Float r, g, b; sc.toLinearRGB(r, g, b);
                t.ref.scale[0] *= r; t.ref.scale[1] *= g; t.ref.scale[2] *= b;
                t.min *= sc; t.max *= sc; t.avg *= sc;
            } else if (cls == "BitmapTexture") {         /* src/textures/bitmap.cpp:401-430 */
                Point2 uvOffset(s); Vector2 uvScale(s);
                std::string filename = s->readString();
                uint32_t filterType = s->readUInt(), wrapU = s->readUInt(), wrapV = s->readUInt();
                Float gamma = s->readFloat();
                Float maxAnisotropy = s->readFloat();
                std::string channel = s->readString();
                size_t size = s->readSize();
                if (!channel.empty()) err("bitmap `channel` selection is not supported by cudapath");
                /* the image file itself: not decoded again (see below), and mostly not even stored (FieldStream) */
                const FieldStream *fs = dynamic_cast<const FieldStream *>(s);
                const size_t dropped = fs ? fs->dropped(s->getPos()) : 0;
                if (dropped > size) err("bitmap texture: inconsistent serialization stream");
                s->seek(s->getPos() + (size - dropped));
                if (filterType > 3) err("unknown bitmap filter type");
                /* One upload per live texture object, however many BSDFs (and serializations) refer to it */
                const SerializableObject *obj = m_idToObj.count(id) ? m_idToObj[id] : NULL;
                if (!obj) err("bitmap texture: object lookup failed");
                std::map<const SerializableObject *, TexInfo>::const_iterator hit = m_texByObject.find(obj);
                if (hit != m_texByObject.end()) { seen[id] = hit->second; return hit->second; }
                /* Level 0 of the texture's own MIP map (TMIPMap::toBitmap, mipmap.h:486-497): the texels exactly as the
                   reference looks them up -- gamma-corrected, clamped, rounded to half */
                /* Default (exactTextures): decode the image file and hand over the unrounded floats -- the library then builds
                   the same pyramid as the reference, bit for bit (about 0.05 s per 4096^2 PFM image, more for compressed
                   formats).  Fast route: level 0 of the texture's own MIP map as half texels; identical at level 0, but the
                   coarser levels are then resampled from half-rounded texels where the reference resamples the unrounded
                   image: a fraction of the texels of levels >= 1 differ by one half ulp (4.9e-4 relative). */
                const Texture *texObj = static_cast<const Texture *>(obj);
                const double tG = now();
                ref<Bitmap> level0;
                if (m_owner->m_exactTextures && !filename.empty() && fs::exists(fs::path(filename))) {
                    ref<FileStream> is = new FileStream(fs::path(filename), FileStream::EReadOnly);
                    ref<Bitmap> bmp = new Bitmap(Bitmap::EAuto, is);
                    if (gamma != 0) bmp->setGamma(gamma);
                    const bool monoFile = bmp->getPixelFormat() == Bitmap::ELuminance || bmp->getPixelFormat() == Bitmap::ELuminanceAlpha;
                    level0 = bmp->convert(monoFile ? Bitmap::ELuminance : Bitmap::ERGB, Bitmap::EFloat32, 1.0f);
                } else
                    level0 = texObj->getBitmap(Vector2i(-1, -1));
                m_tGetBitmap += now() - tG;
                const bool mono = level0->getPixelFormat() == Bitmap::ELuminance || level0->getPixelFormat() == Bitmap::ELuminanceAlpha;
                (void) gamma;
                MtsTextureDesc d;
                memset(&d, 0, sizeof(d));
                d.width = level0->getWidth(); d.height = level0->getHeight(); d.channels = mono ? 1 : 3;
                /* ReconstructionFilter::EBoundaryCondition: EClamp 0, ERepeat 1, EMirror 2, EZero 3, EOne 4 */
                static const int wrapMap[5] = { MTS_WRAP_CLAMP, MTS_WRAP_REPEAT, MTS_WRAP_MIRROR, MTS_WRAP_ZERO, MTS_WRAP_ONE };
                d.wrap_u = wrapMap[std::min(wrapU, 4u)]; d.wrap_v = wrapMap[std::min(wrapV, 4u)];
                d.filter = (int32_t) filterType;            /* EMIPFilterType: ENearest 0, EBilinear 1, ETrilinear 2, EEWA 3 = MTS_FILTER_* */
                d.max_anisotropy = maxAnisotropy;           /* already forced to 1 for non-EWA filters (bitmap.cpp:232-233) */
                d.uv_scale[0] = uvScale.x; d.uv_scale[1] = uvScale.y; d.uv_offset[0] = uvOffset.x; d.uv_offset[1] = uvOffset.y;
                /* the half texels go to the library as they are (MTS_TEXEL_F16); one parallel pass over them for the statistics
                   (Texture::getMinimum / getMaximum / getAverage) */
                const size_t n = (size_t) d.width * d.height;
                const bool isHalf = level0->getComponentFormat() == Bitmap::EFloat16;
                if ((!isHalf && level0->getComponentFormat() != Bitmap::EFloat32) || level0->getChannelCount() != d.channels)
                    err("bitmap texture: unexpected texel storage format");
                const double tC = now();
                const half *src = (const half *) level0->getData();
                const float *srcF = (const float *) level0->getData();
                d.texel_format = isHalf ? MTS_TEXEL_F16 : MTS_TEXEL_F32;
                double sum[3] = {0, 0, 0}; float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
                const int ch = d.channels;
                #pragma omp parallel
                {
                    double lsum[3] = {0, 0, 0}; float lmn[3] = {1e30f, 1e30f, 1e30f}, lmx[3] = {-1e30f, -1e30f, -1e30f};
                    #pragma omp for schedule(static) nowait
                    for (long long i = 0; i < (long long) n; ++i) for (int c = 0; c < ch; ++c) {
                        /* statistics on the half values the MIP map stores, whichever form is uploaded */
                        const float v = isHalf ? (float) src[(size_t) i * ch + c] : (float) half(srcF[(size_t) i * ch + c]);
                        lsum[c] += v; lmn[c] = std::min(lmn[c], v); lmx[c] = std::max(lmx[c], v);
                    }
                    #pragma omp critical
                    for (int c = 0; c < ch; ++c) { sum[c] += lsum[c]; mn[c] = std::min(mn[c], lmn[c]); mx[c] = std::max(mx[c], lmx[c]); }
                }
                if (ch == 1) { sum[1] = sum[2] = sum[0]; mn[1] = mn[2] = mn[0]; mx[1] = mx[2] = mx[0]; }
                const double tA = now();
                m_tConvert += tA - tC;
                int tid = mtscuda_add_texture(m_ctx, &d, srcF);
                m_owner->check(m_ctx, tid);
                m_tAddTexture += now() - tA;
                t.ref.scale[0] = t.ref.scale[1] = t.ref.scale[2] = 1; t.ref.tex = tid; t.constant = false;
                t.min.fromLinearRGB(mn[0], mn[1], mn[2]); t.max.fromLinearRGB(mx[0], mx[1], mx[2]);
                t.avg.fromLinearRGB((Float) (sum[0] / n), (Float) (sum[1] / n), (Float) (sum[2] / n));
                m_texByObject[obj] = t;
            } else
                err("texture class " + cls + " is out of scope (bitmap, scale, constants)");
            seen[id] = t;
            return t;
        }

        /* BSDF::ensureEnergyConservation (src/librender/bsdf.cpp:85-113) */
        static TexInfo conserve(TexInfo t, Float maxAllowed = 1.0f) {
            Float actual = t.max.max();
            if (actual > maxAllowed) {
                Float sc = 0.99f * (maxAllowed / actual);
                for (int i = 0; i < 3; ++i) t.ref.scale[i] *= sc;
                t.min *= sc; t.max *= sc; t.avg *= sc;
            }
            return t;
        }

        int parseBsdf(Stream *s, std::map<uint32_t, int> &seen, std::map<uint32_t, TexInfo> &tex) {
            uint32_t id = s->readUInt();
            if (id == 0) err("missing nested BSDF");
            if (seen.count(id)) return seen[id];
            std::string cls = s->readString();
            s->readBool();                               /* BSDF::serialize: m_ensureEnergyConservation */
            MtsBsdfNode n;
            memset(&n, 0, sizeof(n));
            n.child[0] = n.child[1] = -1; n.t0.tex = n.t1.tex = n.t2.tex = -1;
            const int idx = (int) m_nodes.size();
            m_nodes.push_back(n);
            seen[id] = idx;
            if (cls == "SmoothDiffuse") {
                TexInfo r = conserve(parseTexture(s, tex));
                n.type = MTS_BSDF_DIFFUSE; n.t0 = r.ref;
                n.lobes = r.max.max() > 0 ? MTS_EDIFFUSE_REFLECTION : 0;
            } else if (cls == "SmoothConductor") {
                TexInfo r = conserve(parseTexture(s, tex));
                Spectrum eta(s), k(s);
                n.type = MTS_BSDF_CONDUCTOR; n.t0 = r.ref; n.lobes = MTS_EDELTA_REFLECTION;
                setRGB(eta, n.eta); setRGB(k, n.k);
            } else if (cls == "RoughConductor") {
                uint32_t type = s->readUInt(); bool sv = s->readBool();
                TexInfo au = parseTexture(s, tex), av = parseTexture(s, tex), r = conserve(parseTexture(s, tex));
                Spectrum eta(s), k(s);
                n.type = MTS_BSDF_ROUGHCONDUCTOR; n.flags = distrFlags(type, sv);
                n.t0 = r.ref; n.t1 = au.ref; n.t2 = av.ref; n.lobes = MTS_EGLOSSY_REFLECTION;
                setRGB(eta, n.eta); setRGB(k, n.k);
            } else if (cls == "RoughDielectric") {   /* roughdielectric.cpp:233-244 */
                uint32_t type = s->readUInt(); bool sv = s->readBool();
                TexInfo au = parseTexture(s, tex), av = parseTexture(s, tex);
                TexInfo r = conserve(parseTexture(s, tex)), t = conserve(parseTexture(s, tex));
                Float eta = s->readFloat();
                n.type = MTS_BSDF_ROUGHDIELECTRIC; n.flags = distrFlags(type, sv); n.eta[0] = eta;
                n.t0 = r.ref; n.t1 = au.ref; n.t2 = av.ref;
                for (int i = 0; i < 3; ++i) n.k[i] = t.ref.scale[i];   /* fourth texture reference: k / lut slots */
                n.lut[0] = t.ref.tex;
                n.lobes = MTS_EGLOSSY_REFLECTION | MTS_EGLOSSY_TRANSMISSION | MTS_EFRONTSIDE | MTS_EBACKSIDE;
            } else if (cls == "SmoothPlastic") {
                Float eta = s->readFloat(); bool nonlinear = s->readBool();
                TexInfo sp = conserve(parseTexture(s, tex)), df = conserve(parseTexture(s, tex));
                n.type = MTS_BSDF_PLASTIC; n.eta[0] = eta; n.t0 = df.ref; n.t1 = sp.ref;
                if (nonlinear) n.flags |= MTS_BF_NONLINEAR;
                n.f[0] = fresnelDiffuseReflectance(1 / eta, false);      /* plastic.cpp:193-194 */
                n.f[1] = fresnelDiffuseReflectance(eta, false);
                Float dAvg = df.avg.getLuminance(), sAvg = sp.avg.getLuminance();
                n.f[2] = sAvg / (dAvg + sAvg);
                n.lobes = MTS_EDELTA_REFLECTION | MTS_EDIFFUSE_REFLECTION;
            } else if (cls == "RoughPlastic") {
                uint32_t type = s->readUInt(); bool sv = s->readBool();
                TexInfo sp = conserve(parseTexture(s, tex)), df = conserve(parseTexture(s, tex)), al = parseTexture(s, tex);
                Float eta = s->readFloat(); bool nonlinear = s->readBool();
                n.type = MTS_BSDF_ROUGHPLASTIC; n.flags = distrFlags(type, sv) | (nonlinear ? MTS_BF_NONLINEAR : 0);
                n.eta[0] = eta; n.t0 = df.ref; n.t1 = sp.ref; n.t2 = al.ref;
                Float dAvg = df.avg.getLuminance(), sAvg = sp.avg.getLuminance();
                n.f[2] = sAvg / (dAvg + sAvg);
                /* roughplastic.cpp:277-298: reduce the transmittance tables to the slices the kernels interpolate */
                ref<RoughTransmittance> ext = new RoughTransmittance((MicrofacetDistribution::EType) type);
                ext->checkEta(eta); ext->checkAlpha(al.min.average()); ext->checkAlpha(al.max.average());
                ref<RoughTransmittance> in = ext->clone();
                ext->setEta(eta); in->setEta(1 / eta);
                n.f[0] = ext->getAlphaMin(); n.f[1] = ext->getAlphaMax();
                n.lut[1] = (int32_t) ext->m_thetaSamples; n.lut[2] = (int32_t) ext->m_alphaSamples;
                if (al.constant) {
                    Float alpha = (al.ref.scale[0] + al.ref.scale[1] + al.ref.scale[2]) * (1.0f / 3.0f);
                    ext->setAlpha(alpha);
                    n.lut[0] = (int32_t) m_pool.size();
                    m_pool.insert(m_pool.end(), ext->m_trans, ext->m_trans + ext->m_thetaSamples);
                    n.f[3] = 1 - in->evalDiffuse(std::max(alpha, (Float) 1e-4f));
                } else {
                    n.flags |= MTS_BF_RT_2D;
                    n.lut[0] = (int32_t) m_pool.size();
                    m_pool.insert(m_pool.end(), ext->m_trans, ext->m_trans + ext->m_alphaSamples * ext->m_thetaSamples);
                    n.lut[3] = (int32_t) m_pool.size();
                    m_pool.insert(m_pool.end(), in->m_diffTrans, in->m_diffTrans + in->m_alphaSamples);
                }
                n.lobes = MTS_EGLOSSY_REFLECTION | MTS_EDIFFUSE_REFLECTION;
            } else if (cls == "SmoothDielectric") {
                Float eta = s->readFloat();
                TexInfo r = conserve(parseTexture(s, tex)), t = conserve(parseTexture(s, tex));
                n.type = MTS_BSDF_DIELECTRIC; n.eta[0] = eta; n.t0 = r.ref; n.t1 = t.ref;
                n.lobes = MTS_EDELTA_REFLECTION | MTS_EDELTA_TRANSMISSION | MTS_EBACKSIDE;
            } else if (cls == "SmoothCoating") {
                Float eta = s->readFloat(), thickness = s->readFloat();
                int nested = parseBsdf(s, seen, tex);
                TexInfo sig = parseTexture(s, tex), sp = conserve(parseTexture(s, tex));
                n.type = MTS_BSDF_COATING; n.eta[0] = eta; n.f[0] = thickness; n.t0 = sig.ref; n.t1 = sp.ref; n.child[0] = nested;
                Float avgAbsorption = (sig.avg * (-2 * thickness)).exp().average();   /* coating.cpp:181-183 */
                n.f[2] = 1.0f / (avgAbsorption + 1.0f);
                n.lobes = m_nodes[nested].lobes | MTS_EDELTA_REFLECTION | MTS_EBACKSIDE;
            } else if (cls == "BlendBSDF") {
                TexInfo w = parseTexture(s, tex);
                int a = parseBsdf(s, seen, tex), b = parseBsdf(s, seen, tex);
                n.type = MTS_BSDF_BLEND; n.t0 = w.ref; n.child[0] = a; n.child[1] = b;
                n.lobes = m_nodes[a].lobes | m_nodes[b].lobes;
            } else if (cls == "BumpMap") {
                int nested = parseBsdf(s, seen, tex);
                TexInfo d = parseTexture(s, tex);
                n.type = MTS_BSDF_BUMPMAP; n.t0 = d.ref; n.child[0] = nested; n.lobes = m_nodes[nested].lobes;
            } else if (cls == "ThinDielectric") {   /* thindielectric.cpp:100-106 */
                Float eta = s->readFloat();
                TexInfo r = conserve(parseTexture(s, tex)), t = conserve(parseTexture(s, tex));
                n.type = MTS_BSDF_THINDIELECTRIC; n.eta[0] = eta; n.t0 = r.ref; n.t1 = t.ref;
                n.lobes = MTS_EDELTA_REFLECTION | MTS_ENULL | MTS_EFRONTSIDE | MTS_EBACKSIDE;
            } else if (cls == "Mask") {   /* mask.cpp:76-80: opacity texture, then the nested BSDF */
                TexInfo op = conserve(parseTexture(s, tex));
                int nested = parseBsdf(s, seen, tex);
                n.type = MTS_BSDF_MASK; n.t0 = op.ref; n.child[0] = nested;
                n.lobes = m_nodes[nested].lobes | MTS_ENULL | MTS_EFRONTSIDE | MTS_EBACKSIDE;
            } else if (cls == "TwoSidedBRDF") {   /* twosided.cpp:77-83: BSDF::serialize, then the two nested instances */
                int a = parseBsdf(s, seen, tex), b = parseBsdf(s, seen, tex);
                n.type = MTS_BSDF_TWOSIDED; n.child[0] = a; n.child[1] = b;
                n.lobes = ((m_nodes[a].lobes & ~MTS_EBACKSIDE) | MTS_EFRONTSIDE) | ((m_nodes[b].lobes & ~MTS_EFRONTSIDE) | MTS_EBACKSIDE);
            } else if (cls == "MixtureBSDF") {   /* mixturebsdf.cpp:105-113: count, then (weight, nested instance) pairs */
                size_t count = s->readSize();
                if (count < 1 || count > MTS_MIXTURE_MAX) err("mixturebsdf: 1..16 nested BSDFs are supported");
                std::vector<float> weights(count), cdf(count + 1, 0.0f);
                std::vector<int> kids(count);
                for (size_t i = 0; i < count; ++i) { weights[i] = s->readFloat(); kids[i] = parseBsdf(s, seen, tex); }
                /* the weights are the configured ones (already rescaled, mixturebsdf.cpp:131-143); m_pdf as configure() builds
                   it: DiscreteDistribution::append / normalize (include/mitsuba/core/pmf.h:58-66, 101-114) */
                for (size_t i = 0; i < count; ++i) cdf[i + 1] = cdf[i] + weights[i];
                if (!(cdf[count] > 0)) err("The weights must sum to a value greater than zero!");
                const float norm = 1.0f / cdf[count];
                for (size_t i = 1; i <= count; ++i) cdf[i] *= norm;
                cdf[count] = 1.0f;
                n.type = MTS_BSDF_MIXTURE; n.lut[0] = (int32_t) m_pool.size(); n.lut[1] = (int32_t) count;
                for (size_t i = 0; i < count; ++i) { m_pool.push_back((float) kids[i]); n.lobes |= m_nodes[kids[i]].lobes; }
                m_pool.insert(m_pool.end(), weights.begin(), weights.end());
                m_pool.insert(m_pool.end(), cdf.begin(), cdf.end());
            } else if (cls == "RoughCoating") {   /* roughcoating.cpp:167-178 */
                uint32_t type = s->readUInt(); bool sv = s->readBool();
                int nested = parseBsdf(s, seen, tex);
                TexInfo sig = parseTexture(s, tex), sp = conserve(parseTexture(s, tex)), al = parseTexture(s, tex);
                Float eta = s->readFloat(), thickness = s->readFloat();
                n.type = MTS_BSDF_ROUGHCOATING; n.flags = distrFlags(type, sv); n.eta[0] = eta; n.child[0] = nested;
                n.t0 = sig.ref; n.t1 = sp.ref; n.t2 = al.ref; n.f[3] = thickness;
                Float avgAbsorption = (sig.avg * (-2 * thickness)).exp().average();   /* roughcoating.cpp:200-203 */
                n.f[2] = 1.0f / (avgAbsorption + 1.0f);
                ref<RoughTransmittance> rt = new RoughTransmittance((MicrofacetDistribution::EType) type);   /* :209-222 */
                rt->checkEta(eta); rt->checkAlpha(al.min.average()); rt->checkAlpha(al.max.average());
                rt->setEta(eta);
                n.f[0] = rt->getAlphaMin(); n.f[1] = rt->getAlphaMax();
                n.lut[1] = (int32_t) rt->m_thetaSamples; n.lut[2] = (int32_t) rt->m_alphaSamples;
                n.lut[0] = (int32_t) m_pool.size();
                if (al.constant) {
                    rt->setAlpha((al.ref.scale[0] + al.ref.scale[1] + al.ref.scale[2]) * (1.0f / 3.0f));
                    m_pool.insert(m_pool.end(), rt->m_trans, rt->m_trans + rt->m_thetaSamples);
                } else {
                    n.flags |= MTS_BF_RT_2D;
                    m_pool.insert(m_pool.end(), rt->m_trans, rt->m_trans + rt->m_alphaSamples * rt->m_thetaSamples);
                }
                n.lobes = m_nodes[nested].lobes | MTS_EGLOSSY_REFLECTION | MTS_EFRONTSIDE | MTS_EBACKSIDE;
            } else if (cls == "NormalMap") {   /* normalmap.cpp:74-79: nested BSDF, then the normal texture */
                int nested = parseBsdf(s, seen, tex);
                TexInfo nm = parseTexture(s, tex);
                n.type = MTS_BSDF_NORMALMAP; n.t0 = nm.ref; n.child[0] = nested; n.lobes = m_nodes[nested].lobes;
            } else
                err("BSDF class " + cls + " is out of scope (diffuse, conductor, roughconductor, plastic, roughplastic, "
                    "dielectric, coating, blendbsdf, bumpmap, twosided, thindielectric, mask, roughdielectric, mixturebsdf, "
                    "roughcoating, normalmap)");
            m_nodes[idx] = n;
            return idx;
        }

        static void setRGB(const Spectrum &v, float *out) { Float r, g, b; v.toLinearRGB(r, g, b); out[0] = r; out[1] = g; out[2] = b; }
        int distrFlags(uint32_t type, bool sampleVisible) const {
            if (type != MicrofacetDistribution::EBeckmann && type != MicrofacetDistribution::EGGX)
                err("microfacet distribution is out of scope (beckmann, ggx)");
            return (type == MicrofacetDistribution::EGGX ? MTS_BF_GGX : 0) | (sampleVisible ? MTS_BF_SAMPLE_VISIBLE : 0);
        }

        const CudaPath *m_owner;
        MtsContext *m_ctx;
        std::vector<MtsBsdfNode> m_nodes;
        std::vector<float> m_pool;
        std::map<const BSDF *, int> m_bsdfIds;
    public:
        double m_tSerialize, m_tGetBitmap, m_tConvert, m_tAddTexture, m_tMesh;   /* MTS_PLUGIN_TIMING */
    private:
        std::map<uint32_t, const SerializableObject *> m_idToObj;           /* of the serialization being parsed */
        std::map<const SerializableObject *, TexInfo> m_texByObject;        /* uploaded bitmap textures */
    };

    int m_maxDepth, m_rrDepth, m_device, m_devices;
    bool m_strictNormals, m_hideEmitters, m_exactTextures;
    uint32_t m_seed;
    volatile int m_cancel;
};

MTS_IMPLEMENT_CLASS_S(CudaPath, false, Integrator)
MTS_EXPORT_PLUGIN(CudaPath, "B200-native CUDA path tracer");
MTS_NAMESPACE_END
