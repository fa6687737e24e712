/* qaray_b200.h — C ABI of the B200 renderer for qaray's per-pixel Monte-Carlo loop.
 *
 * The reference (wilsonCernWq/qaray) has no FFI: its plug-in seam is the C++ `Renderer` subclass contract
 *     r->Init(); LoadScene(file); r->ComputeScene(renderImage, scene); r->Render(); r->Terminate();
 * (src/main.cpp:55-59, src/renderers/renderer.h:101-110), with `Renderer::ThreadRender` -> `PixelRender`
 * (src/renderers/renderer.cpp:302-423) as the hot path. This header is what a binding for that seam talks to:
 * plain pointers and sizes, one opaque scene handle, no C++ or torch types. Every entry point names the
 * reference interface it replaces. All functions return 0 on success or a negative QG_E* code; the message is
 * available from qg_last_error() (thread-local). The library copies every host buffer it is given.
 *
 * There is no CPU fallback: every compute entry point fails with QG_ENODEVICE when no CUDA device is usable.
 */
#ifndef QARAY_B200_H
#define QARAY_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QG_OK 0
#define QG_EINVAL (-1)   /* bad argument */
#define QG_EIO (-2)      /* scene / asset file could not be read (LoadScene returning 0, xmlload.cpp:71-95) */
#define QG_ENODEVICE (-3) /* no usable CUDA device */
#define QG_ECUDA (-4)    /* a CUDA call failed */
#define QG_ENOMEM (-5)

/* object kinds: Sphere / Plane / TriObj (src/objects/objects.h) */
enum { QG_OBJ_NONE = 0, QG_OBJ_SPHERE = 1, QG_OBJ_PLANE = 2, QG_OBJ_MESH = 3 };
/* light kinds (src/lights/lights.h:35-171) */
enum { QG_LIGHT_AMBIENT = 0, QG_LIGHT_DIRECT = 1, QG_LIGHT_POINT = 2, QG_LIGHT_SPOT = 3 };
/* texture kinds (src/textures/texture.h): NULLTEX = a TextureMap whose file failed to load (samples black) */
enum { QG_TEX_NULL = 0, QG_TEX_CHECKER = 1, QG_TEX_IMAGE = 2 };
/* material kinds: MtlBlinn (alias of MtlBlinn_PhotonMap, src/materials/materials.h:57) and MultiMtl (:65-87) */
enum { QG_MTL_BLINN = 0, QG_MTL_MULTI = 1 };

/* the classes `MtlBlinn` can alias (src/materials/materials.h:57-61): MtlBlinn_PhotonMap (.cpp:256-500, photon maps off),
 * MtlBlinn_PathTracing (.cpp:69-300), MtlBlinn_MonteCarloGI (.cpp:34-266), MtlBlinn_Basic (.cpp:30-185), MtlPhong_Basic */
enum { QG_CLASS_PHOTONMAP = 0, QG_CLASS_PATHTRACING = 1, QG_CLASS_MONTECARLOGI = 2, QG_CLASS_BLINN_BASIC = 3, QG_CLASS_PHONG_BASIC = 4 };

/* Transformation (src/core/transform.h:36-79). Matrices column-major like glm::mat3. */
typedef struct qgXform {
  float tm[9];
  float itm[9];
  float pos[3];
} qgXform;

/* Node of the scene tree (src/core/node.h), listed in pre-order = the order Scene::TraceNodeNormal visits. */
typedef struct qgNode {
  int32_t parent;      /* -1 for the root (node 0, no object, identity transform) */
  int32_t object_type; /* QG_OBJ_* */
  int32_t mesh;        /* index into meshes[] when object_type == QG_OBJ_MESH */
  int32_t material;    /* index into materials[] or -1 */
  qgXform xf;
} qgNode;

/* TriMesh after TriObj::Load (src/objects/objects.h:65-73): faces already in the reference's sorted order. */
typedef struct qgMesh {
  const float *vertices;       /* 3 * num_vertices */
  const float *normals;        /* 3 * num_normals */
  const float *texcoords;      /* 2 * num_texcoords (may be NULL when 0) */
  const int32_t *face_indices; /* 9 per face: v0 n0 t0 v1 n1 t1 v2 n2 t2; t = -1 when the corner has no uv */
  const int32_t *face_material; /* per face material id (MultiMtl sub-material), -1 = none */
  int32_t num_vertices, num_normals, num_texcoords, num_faces;
  float bound_min[3], bound_max[3]; /* TriMesh::ComputeBoundingBox (TriMesh.cpp:117-133) */
} qgMesh;

/* TexturedColor + TextureMap (src/core/texture.h:55-96) */
typedef struct qgTexturedColor {
  float color[3];
  int32_t texture; /* -1: no map; else index into textures[] */
  qgXform xf;      /* the TextureMap's transformation */
} qgTexturedColor;

typedef struct qgTexture {
  int32_t type; /* QG_TEX_* */
  float color1[3], color2[3]; /* checker */
  int32_t width, height;
  const uint8_t *rgb8; /* image: width*height*3, row 0 first */
} qgTexture;

typedef struct qgMaterial {
  int32_t kind; /* QG_MTL_* */
  qgTexturedColor diffuse, specular, emission, reflection, refraction;
  float absorption[3];
  float ior;
  float glossiness;            /* specular exponent */
  float reflection_glossiness; /* as stored by SetReflectionGlossiness: value or -1 (MtlBlinn_PhotonMap.cpp:56-63) */
  float refraction_glossiness;
  float kill;                  /* russian-roulette absorb weight, 0.1 */
  int32_t sub_first, sub_count; /* QG_MTL_MULTI: materials[multi[sub_first + mtlID]] */
} qgMaterial;

typedef struct qgLight {
  int32_t type; /* QG_LIGHT_* */
  float intensity[3];
  float position[3];
  float direction[3]; /* direct: normalised direction; spot: axis */
  float size;
  float inner, outer; /* spot cone in tan space (lights.cpp:121-128) */
} qgLight;

/* Camera after LoadScene (dir normalised, up re-orthogonalised; xmlload.cpp:116-144) */
typedef struct qgCamera {
  float pos[3], dir[3], up[3];
  float fovy, focal, dof;
  int32_t width, height;
} qgCamera;

typedef struct qgSceneDesc {
  const qgNode *nodes;
  const qgMesh *meshes;
  const qgMaterial *materials;
  const int32_t *multi; /* sub-material index lists of the MultiMtl entries */
  const qgLight *lights;
  const qgTexture *textures;
  int32_t num_nodes, num_meshes, num_materials, num_multi, num_lights, num_textures;
  qgTexturedColor background, environment;
  qgCamera camera;
} qgSceneDesc;

/* DiffRay (src/core/ray.h:36-72): centre ray plus the two differential rays; has_diff = 0 for secondary rays. */
typedef struct qgRay {
  float p[3], dir[3];
  float xp[3], xdir[3];
  float yp[3], ydir[3];
  int32_t has_diff;
} qgRay;

/* What Scene::TraceNodeNormal leaves in DiffHitInfo::c (src/core/hitinfo.h:36-52), plus the primitive id. */
typedef struct qgHit {
  int32_t hit;   /* 0/1 */
  int32_t node;  /* pre-order index of hInfo.c.node, -1 on miss */
  int32_t face;  /* face id inside the mesh (the reference's faceID into TriMesh::F), -1 for sphere/plane */
  int32_t mtl_id;
  int32_t front;
  int32_t has_texture;
  float z;
  float p[3], N[3], uvw[3], duvw[6];
} qgHit;

typedef struct qgRenderParams {
  int32_t spp;         /* RendererParam::sppMax (src/renderers/renderer.h:47-68); `-spp N` sets min = max (main.cpp:21-24) */
  int32_t max_bounce;  /* Material::maxBounce (src/core/material.cpp:31), default 5 */
  int32_t srgb;        /* RendererParam::useSRGB */
  uint32_t seed;       /* stream seed of the counter-based RNG that replaces src/samplers */
  int32_t width, height; /* 0 = the scene camera's; overrides <width>/<height> like editing the XML */
  int32_t tile_rank, tile_ranks; /* Renderer::mpiRank / mpiSize: this shard renders 4x4 tiles k = rank (mod ranks) (renderer.cpp:383-387) */
  int32_t sample_begin, sample_end; /* sample-index shard [begin,end); 0,0 = [0,spp) */
  int32_t batch_paths; /* paths in flight per wave; 0 = default */
  int32_t sort_by_material; /* 1 = shade every depth's hits in material order (counting sort of the shade queue); same
                               image, different warp composition. Default 0: measured slower on the BASELINE scenes */
  int32_t material_class; /* QG_CLASS_*: which MtlBlinn class shades `blinn` materials (the reference picks it with a
                             compile-time alias, src/materials/materials.h:57-61); 0 = MtlBlinn_PhotonMap, its default */
  int32_t spp_min; /* RendererParam::sppMin when below spp (= sppMax): SuperSamplerHalton's adaptive stop
                      (src/scene/scene.cpp:92-97); 0 or spp = fixed sample count */
  int32_t use_photon_map; /* Scene::usePhotonMap (`-use-photon-map`, src/main.cpp:35): diffuse hits gather from the maps
                             qg_photon_maps_build made (MtlBlinn_PhotonMap.cpp:341-352,426-458); material_class 0 only */
} qgRenderParams;

typedef struct qgRenderStats {
  uint64_t samples;       /* PixelRender loop iterations */
  uint64_t closest_rays;  /* Scene::TraceNodeNormal calls on the root */
  uint64_t shadow_rays;   /* Scene::TraceNodeShadow calls on the root */
  uint64_t kernel_launches;
  uint64_t nodes_visited, tris_tested, objects_tested; /* closest-hit rays; filled by the counting build only, else 0 */
  uint64_t shadow_nodes_visited, shadow_tris_tested, shadow_objects_tested; /* same for any-hit rays */
  double device_ms;       /* CUDA-event time of the render on its stream */
  double trace_ms, shade_ms, shadow_ms; /* device time per kernel class: closest-hit (+ hit attributes), shade, and the any-hit
                                           time not already inside a closest-hit interval (the any-hit kernels of depth d run
                                           beside the closest-hit kernel of depth d + 1): trace_ms + shadow_ms = time some
                                           traversal kernel was running */
} qgRenderStats;

typedef struct qgSceneInfo {
  int32_t num_nodes, num_meshes, num_materials, num_lights, num_textures;
  int32_t num_object_nodes;
  int64_t num_faces, num_bvh_nodes;
  int64_t device_bytes;
  qgCamera camera;
  float screen_u[3], screen_v[3], screen_a[3]; /* Renderer::ComputeScene, renderer.cpp:77-91 */
} qgSceneInfo;

typedef struct qgScene qgScene;

/* Renderer::Init (src/renderers/renderer.h:101). device = CUDA ordinal. */
int qg_init(int device);
/* Renderer::Terminate */
void qg_shutdown(void);
const char *qg_last_error(void);

/* LoadScene(filename) (src/parser/xmlload.cpp:71) + Renderer::ComputeScene's upload. asset_dir: directory the
 * XML's relative .obj/.png names resolve against (the reference resolves them against the process cwd). */
int qg_scene_load_xml(const char *xml_path, const char *asset_dir, qgScene **out);
/* The host half of LoadScene alone (XML + OBJ + PNG -> flattened tables); needs no GPU. A binding can inspect or
 * edit the description and hand it to qg_scene_create. */
typedef struct qgHostScene qgHostScene;
int qg_host_scene_load_xml(const char *xml_path, const char *asset_dir, qgHostScene **out);
const qgSceneDesc *qg_host_scene_desc(const qgHostScene *hs);
const char *qg_host_scene_log(const qgHostScene *hs); /* what the reference would have printed while loading */
void qg_host_scene_destroy(qgHostScene *hs);
/* Same, from an already flattened scene (what a C++ binding fills from qaray::Scene, src/scene/scene.h:55-67). */
int qg_scene_create(const qgSceneDesc *desc, qgScene **out);
/* The BVH built for `mesh`, in the array form of the reference's cyBVH (src/ext/cyBVH.h:63-166: BVH::Build with MeanSplit,
 * ConvertTempData; TriObj::Load builds it with 4 elements per leaf, src/objects/objects.h:71): slot 0 unused, root = 1, sibling
 * nodes adjacent. boxes6 = min xyz, max xyz per slot; data = the node word (leaf: bit 31 | (count - 1) << 28 | element
 * offset; internal: index of the first child); elements = face ids in leaf order. Host only (no device needed) — lets a
 * binding or a test compare the topology with cyBVH's node for node. Any output may be NULL (sizes only). */
int qg_mesh_bvh(const qgMesh *mesh, float *boxes6, uint32_t *data, int64_t node_capacity, int64_t *num_nodes, uint32_t *elements,
                int64_t element_capacity, int64_t *num_elements);
void qg_scene_destroy(qgScene *scene);
int qg_scene_info(const qgScene *scene, qgSceneInfo *out);
/* copies the flattened node table (num_nodes entries) — lets a binding check its transforms */
int qg_scene_nodes(const qgScene *scene, qgNode *out, int32_t capacity);

/* Scene::TraceNodeNormal(rootNode, ray, hInfo) over a ray buffer (src/scene/scene.cpp:50-74). Host buffers. */
int qg_trace_closest(qgScene *scene, const qgRay *rays, size_t n, qgHit *out);
/* Scene::TraceNodeShadow(rootNode, ray, hInfo with z = tmax) (src/scene/scene.cpp:35-46). rays6 = p, dir. */
int qg_trace_any(qgScene *scene, const float *rays6, const float *tmax, size_t n, uint8_t *occluded);
/* qg_trace_any without an allocation. NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * d_rays6 (6n floats), d_tmax (n floats) and d_occluded (n bytes) are DEVICE pointers on the scene's device. The launch's
 * scratch (work cursor + the hand-over queue of the exact walk) is a static device array of 4 MiB + 256 B that every GPU
 * loading the library holds until exit, so n <= QG_NOALLOC_MAX_RAYS; a larger n is refused with QG_EINVAL. The call makes no
 * cudaMalloc / cudaFree and does not synchronise: work is enqueued on `stream` (a cudaStream_t; NULL = the scene's own
 * stream) and d_occluded is complete once that stream has passed this point. All calls on one device share the scratch:
 * calls that overlap on the GPU (different streams or threads) race — serialise them on one stream. Results are those of
 * qg_trace_any, bit for bit. */
#define QG_NOALLOC_MAX_RAYS (1 << 20)
int qg_trace_any_noalloc(qgScene *scene, const float *d_rays6, const float *d_tmax, size_t n, uint8_t *d_occluded, void *stream);
/* Primary DiffRay of pixel (i,j), sample s as PixelRender builds it (renderer.cpp:312-326). pix = i,j,s triples. */
int qg_primary_rays(qgScene *scene, const int32_t *pix, size_t n, int32_t width, int32_t height, qgRay *out);
/* One PixelRender loop body per (i,j,s) triple: the sample's radiance before accumulation (renderer.cpp:312-339). */
int qg_sample_radiance(qgScene *scene, const qgRenderParams *params, const int32_t *pix, size_t n, float *rgb, float *z);

/* Renderer::Render for the batch renderer: ThreadRender + the FrameBuffer planes PixelRender fills
 * (renderer.cpp:348-365). Host outputs, any may be NULL: rgb_mean W*H*3 linear means, rgb8 W*H*3,
 * depth W*H, sample_count W*H (255*s/sppMax), mask W*H. */
int qg_render(qgScene *scene, const qgRenderParams *params, float *rgb_mean, uint8_t *rgb8, float *depth,
              uint8_t *sample_count, uint8_t *mask, qgRenderStats *stats);
/* Shard render into a DEVICE accumulation buffer accum[W*H*4] = {sum R, sum G, sum B, sum weight} (+=) and
 * depth[W*H] (written where this shard owns sample 0, else left untouched). For the multi-GPU path: every rank
 * accumulates its tile/sample shard, the buffers are summed by one NCCL reduce (replacing the MPI_Send/Recv
 * gather of src/renderers/Renderer_MPI.cpp:142-207), and the root resolves. stream = cudaStream_t, or NULL for the scene's own non-blocking
 * stream (NOT ordered with the legacy default stream: synchronise d_accum / d_depth producers first). The any-hit kernels
 * of each depth run on a second, library-owned stream that forks from and joins `stream` inside the call; the call returns
 * after the whole render has finished on the device. */
int qg_render_accumulate(qgScene *scene, const qgRenderParams *params, float *d_accum, float *d_depth, void *stream,
                         qgRenderStats *stats);
/* accum (device) -> mean, sRGB, clamp, round (renderer.cpp:348-361). Host outputs as in qg_render. */
int qg_resolve(qgScene *scene, const qgRenderParams *params, const float *d_accum, const float *d_depth,
               float *rgb_mean, uint8_t *rgb8, float *depth, uint8_t *sample_count, uint8_t *mask, void *stream);

/* Host-only (needs no GPU): how many samples of each pixel the shard described by params renders — the tile loop of
 * Renderer::ThreadRender (k = mpiRank; k += mpiSize over 4x4 tiles, renderer.cpp:383-398) restricted to
 * [sample_begin, sample_end). counts[width*height]. The kernels enumerate paths with the same code, so the shards of
 * all ranks must tile the frame exactly once per sample (tests/test_sharding.py). */
int qg_shard_coverage(const qgCamera *camera, const qgRenderParams *params, uint16_t *counts);
/* Host-only: the sizes the render driver derives for that shard, in 64 bits (3840 x 2160 at 1024 spp is 8.5e9 samples):
 * pixels the shard owns, samples = pixels x its sample range (PixelRender loop iterations, renderer.cpp:310), and path
 * slots enumerated (16 per 4x4 tile, edge tiles padded). Any output may be NULL. */
int qg_shard_size(const qgCamera *camera, const qgRenderParams *params, uint64_t *pixels, uint64_t *samples, uint64_t *paths);

/* ---- photon / caustics maps: the emission pass of Renderer::ComputeScene (src/renderers/renderer.cpp:118-291) -------------
 * Photons leave the point lights (the only IsPhotonSource, src/lights/lights.h:114), are traced with Scene::TraceNodeNormal,
 * stored at diffuse surfaces from their second segment on, and continued by MtlBlinn_PhotonMap::RandomPhotonBounce
 * (src/materials/MtlBlinn_PhotonMap.cpp:503-575) until the map holds `size` photons; powers are then divided by the number
 * of emissions that stored something. Defaults = RendererParam's (src/renderers/renderer.h:52-57). */
typedef struct qgPhotonParams {
  int32_t photon_map_size, photon_map_bounce;     /* 10000, 20 */
  float photon_map_radius;                        /* 0.2 */
  int32_t caustics_map_size, caustics_map_bounce; /* 1000, 20 */
  float caustics_map_radius;                      /* 1.0 */
  uint32_t seed;                                  /* stream seed of the counter-based RNG (emission e has its own stream) */
} qgPhotonParams;
typedef struct qgPhotonStats {
  uint64_t photons[2];   /* stored: [0] photon map, [1] caustics map */
  uint64_t emitted[2];   /* numOfEmittedRays: emissions that stored at least one photon (the power normalisation) */
  uint64_t emissions[2]; /* emissions traced */
  uint64_t rays[2];      /* Scene::TraceNodeNormal calls */
  double device_ms;
} qgPhotonStats;
/* builds both maps for the scene and keeps them with it (Renderer::ComputeScene); a render with use_photon_map = 1 needs them.
 * A scene without point lights, or a size of 0, leaves the maps empty (the reference skips the pass, renderer.cpp:120-123). */
int qg_photon_maps_build(qgScene *scene, const qgPhotonParams *params, qgPhotonStats *stats);
/* Renderer::Terminate: scene->photonmap.Clear(); scene->causticsmap.Clear() */
void qg_photon_maps_clear(qgScene *scene);
/* The photons of map `which` (0 photon map, 1 caustics map) as the estimate reads them — position, direction and power after
 * cyPhotonMap's quantisation (src/ext/cyPhotonMap.h:166-206) and ScalePhotonPowers — in emission order. Host arrays of
 * 3 * capacity floats each (any may be NULL); *count = photons in the map. */
int qg_photon_map_read(qgScene *scene, int32_t which, float *pos3, float *dir3, float *power3, size_t capacity, size_t *count);
/* The body of the emission loop for emissions [first, first + n) of map `which`, without the map-full stop: counts[e] photons
 * (at most bounce each) in records[(e * bounce + k) * 9] = position, incoming direction, power BEFORE the normalisation. */
int qg_photon_emit(qgScene *scene, const qgPhotonParams *params, int32_t which, uint32_t first, size_t n, int32_t *counts, float *records);
/* cyPhotonMap::EstimateIrradiance<100>(I, D, radius, pos, &normal, 1, FILTER_TYPE_QUADRATIC) (cyPhotonMap.h:333-395) at n points */
int qg_photon_estimate(qgScene *scene, int32_t which, const float *pos3, const float *normal3, size_t n, float *irrad3, float *dir3);
/* replaces map `which` by caller-supplied photons: AddPhoton (position, direction, power; quantised as cyPhotonMap does) and
 * ScalePhotonPowers(power_scale) — lets a binding or a test hand over a map built elsewhere */
int qg_photon_map_set(qgScene *scene, int32_t which, const float *pos3, const float *dir3, const float *power3, size_t n, float radius,
                      float power_scale);

/* ---- multi-GPU: the reference's rank model (Renderer_MPI, src/renderers/Renderer_MPI.cpp) over NCCL ----------------------
 * Every rank accumulates its shard of the frame — tiles k = r_t (mod G_t) of the reference's interleaved 4x4-tile rule
 * (renderer.cpp:383-387) and the r_s-th of G_s equal sample ranges, G = G_t x G_s ranks, r = r_s * G_t + r_t — into a full-frame
 * accumulation buffer; one ncclReduce(sum) per buffer to rank 0 replaces the MPI_Send / MPI_Recv / PlaceImage gather
 * (Renderer_MPI.cpp:142-207). NCCL is bound when the first of these is called (dlopen libnccl.so.2); without it they fail
 * with QG_ENODEVICE and everything above keeps working. */

/* this rank's shard of `params` (which must describe the whole frame). tile_split = G_t: 1 = sample shards only,
 * nranks = tile shards only (the reference's own rule). Needs no GPU. */
int qg_shard_params(const qgRenderParams *params, int32_t rank, int32_t nranks, int32_t tile_split, qgRenderParams *out);

/* One process per GPU (Renderer_MPI::Init, :33-55: MPI_Comm_rank / MPI_Comm_size). Rank 0 fills id[QG_COMM_ID_BYTES] and ships
 * it to the other ranks by any means; every rank then joins on the device qg_init chose. */
#define QG_COMM_ID_BYTES 128
typedef struct qgComm qgComm;
int qg_comm_unique_id(void *id);
int qg_comm_create(const void *id, int rank, int nranks, qgComm **out);
int qg_comm_info(const qgComm *comm, int32_t *rank, int32_t *nranks, int32_t *nccl_version);
void qg_comm_destroy(qgComm *comm);
/* Renderer_MPI::Render (:123-207) up to the resolve: this rank's shard of `params` into d_accum / d_depth (device, zeroed by
 * the caller, as for qg_render_accumulate), then the reduce to rank 0 on the same stream; returns when both have finished.
 * Rank 0 then calls qg_resolve. stats = this rank's. */
int qg_comm_render(qgScene *scene, qgComm *comm, const qgRenderParams *params, int32_t tile_split, float *d_accum, float *d_depth,
                   void *stream, qgRenderStats *stats);

/* One process, several GPUs: ncclCommInitAll over `devices` (NULL = 0..ndev-1), a scene replica, a host thread and a frame
 * buffer per device. qg_multi_render = qg_render over all of them: host outputs as for qg_render, stats summed over the ranks
 * (times = the slowest rank's). rank_image_prefix != NULL also writes <prefix>rank_<r>_{color,depth,sample}Buffer.png, the
 * per-rank planes Renderer_MPI::Render saves before its gather (:132-138). */
typedef struct qgMulti qgMulti;
int qg_multi_create(const int32_t *devices, int32_t ndev, qgMulti **out);
int qg_multi_load_xml(qgMulti *multi, const char *xml_path, const char *asset_dir);
int qg_multi_create_scene(qgMulti *multi, const qgSceneDesc *desc);
qgScene *qg_multi_scene(qgMulti *multi, int32_t rank); /* rank's replica (owned by multi) */
int qg_multi_render(qgMulti *multi, const qgRenderParams *params, int32_t tile_split, const char *rank_image_prefix, float *rgb_mean,
                    uint8_t *rgb8, float *depth, uint8_t *sample_count, uint8_t *mask, qgRenderStats *stats);
void qg_multi_destroy(qgMulti *multi);

/* FrameBuffer::SavePNG family (src/fb/framebuffer.cpp:62-143): colour, normalised depth, sample-count images. */
int qg_save_png_rgb8(const char *path, const uint8_t *rgb8, int32_t width, int32_t height);
int qg_save_png_depth(const char *path, const float *depth, int32_t width, int32_t height);
int qg_save_png_gray8(const char *path, const uint8_t *gray, int32_t width, int32_t height);
/* FrameBuffer::ComputeSampleCountImage + SaveSampleCountImage (framebuffer.cpp:88-107,137-140): the sample-count plane
 * stretched to 255 * (s - smin) / (smax - smin), all zero when every pixel took the same number of samples */
int qg_save_png_sample_count(const char *path, const uint8_t *sample_count, int32_t width, int32_t height);

#ifdef __cplusplus
}
#endif
#endif /* QARAY_B200_H */
