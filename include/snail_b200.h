/* snail_b200 — C ABI of libsnail_b200.so.
 *
 * B200-native (sm_100a CUDA) replacement for the hot path of nadult/Snail: SAH-BVH traversal, ray/triangle
 * intersection and the Whitted shading loop (shadow + one reflection bounce). The reference has no FFI layer;
 * its seam is the C++ interface that src/render.cpp and the recursive calls in src/scene_trace.cpp use. Every
 * entry point below names the reference interface it replaces (paths relative to the reference checkout).
 *
 * Conventions: handle based; every int-returning call returns 0 on success and a negative snl_status on failure
 * (message through snl_last_error(), thread local); no exceptions cross the boundary; all pointers are caller
 * owned; "host" pointers are ordinary host memory, "dev" pointers are CUDA device memory of the scene's device.
 * There is no CPU fallback: without a CUDA device snl_scene_create fails with SNL_ERR_NO_DEVICE.
 */
#ifndef SNAIL_B200_H
#define SNAIL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SNL_ABI_VERSION 3

typedef enum snl_status {
	SNL_OK = 0,
	SNL_ERR_INVALID = -1,    /* bad argument */
	SNL_ERR_NO_DEVICE = -2,  /* no CUDA device / device index out of range */
	SNL_ERR_CUDA = -3,       /* a CUDA runtime call failed */
	SNL_ERR_STATE = -4,      /* call sequence error (e.g. render before set_geometry) */
	SNL_ERR_IO = -5,         /* file could not be read / parsed */
	SNL_ERR_UNSUPPORTED = -6 /* option the device path does not implement */
} snl_status;

typedef struct snl_scene snl_scene;

/* Camera: src/camera.h:7-14 `struct Camera { float plane_dist; Vec3f pos, right, up, front; }`, passed in the
 * order of its constructor (pos, right, up, front, plane_dist). */
typedef struct snl_camera {
	float pos[3], right[3], up[3], front[3];
	float plane_dist;
} snl_camera;

/* Light: byte-compatible with src/light.h:5-17 `class Light { Vec3f pos, color; float radius, radSq, iRadius; }`. */
typedef struct snl_light {
	float pos[3], color[3];
	float radius, radSq, iRadius;
} snl_light;

/* Material kinds = the shading::Material subclasses MakeMaterials can produce (src/shading/material.cpp:157-183). */
typedef enum snl_material_kind {
	SNL_MAT_SIMPLE = 0,     /* shading::SimpleMaterial<true>   src/shading/simple_material.h:13-30 */
	SNL_MAT_UBER = 1,       /* shading::UberMaterial           src/shading/uber_material.h:12-31   */
	SNL_MAT_TEX = 2,        /* shading::TexMaterial<true>      src/shading/tex_material.h:14-34    */
	SNL_MAT_TRANSPARENT = 3 /* shading::TransparentMaterial<1> src/shading/transparent_material.h:15-41 */
} snl_material_kind;

/* Material::flags (src/shading/material.h:14-19) */
enum { SNL_MATF_TEXCOORDS = 1, SNL_MATF_REFLECTION = 2, SNL_MATF_REFRACTION = 4, SNL_MATF_TRANSPARENCY = 8 };

typedef struct snl_material {
	int kind;          /* snl_material_kind */
	int flags;         /* SNL_MATF_* */
	float color[3];    /* Simple: colour; Uber: diffuse AFTER the x/z swap of uber_material.cpp:11 */
	float specular[3]; /* Uber only */
	float opacity;     /* Uber: dissolveFactor */
	int texture;       /* Tex, Transparent: index of the colour texture in the texture table, else -1 */
	int mask_texture;  /* Transparent: index of the opacity texture (map_d; its first channel is the opacity), else -1 */
} snl_material;

/* RGB8 mip chain laid out as MipmapTexture does (src/mipmap_texture.cpp:298-308): mips stored back to back,
 * mip m is max(w>>m,1) x max(h>>m,1), pitch 3*width, no padding. w and h must be powers of two
 * (src/sampling/point_sampler.cpp:7-8). */
typedef struct snl_texture {
	const unsigned char *data; /* host */
	size_t size;               /* total bytes of the chain */
	int width, height, mips;
} snl_texture;

/* TreeStats analogue (src/tree_stats.h:35-113). `rays` follows TreeStats::TracingRays semantics
 * (src/scene_trace.cpp:117,557): every traced primary ray + every shadow ray actually cast + every reflection
 * ray + the reflection pass's shadow rays. */
typedef struct snl_stats {
	uint64_t rays;          /* = primary + shadow + reflection + reflection_shadow + transparency */
	uint64_t primary;
	uint64_t shadow;
	uint64_t reflection;
	uint64_t reflection_shadow;
	uint64_t transparency;  /* rays continued behind a transparent hit (TraceTransparency, src/scene_trace.cpp:620-634) */
	uint64_t node_visits;   /* only counted when the library is built with SNL_COUNTERS */
	uint64_t tri_tests;     /* idem: 48-byte per-origin triangle records tested */
	uint64_t gtri_tests;    /* idem: 64-byte general triangle records tested (reflection rays) */
	uint64_t transparency_truncated; /* rays whose chain of transparent surfaces was cut after 8 (the reference's
	                                    recursion is unbounded); 0 = the frame is exactly the reference's */
	float kernel_ms;        /* device time of the last call (CUDA events on the scene's stream): its kernels; for snl_render
	                           with a host image also the row-band device->host copies queued behind them */
} snl_stats;

const char *snl_last_error(void);
int snl_abi_version(void);
int snl_device_count(void);

/* Scene<BVH> life cycle (src/scene.h:13-57). One scene owns one CUDA device + one stream. */
snl_scene *snl_scene_create(int device);
void snl_scene_destroy(snl_scene *);

/* Flatten step. Takes the host builder's output exactly as BVH holds it (src/bvh/tree.h:60-72,86-92):
 *   nodes    BVH::Node[nNodes]     32 B  {BBox bbox; int subNode|first; short axis, firstNode | int count}
 *   tris     Triangle[nTris]       64 B  {a, ba, ca, t0, it0, pad, plane}   src/triangle.h:133-135
 *   shTris   ShTriangle[nTris]     64 B  {uv[3], nrm[3], matId|flat<<31}    src/triangle.h:181-230 (may be NULL)
 *   triMatId BVH::materials[i].id  (scene material index, or ~0 = default)  src/bvh/tree.h:74-84
 * and uploads the 64-byte two-child-box node layout + 16-byte aligned triangle records described in DESIGN.md.
 * Float fields are copied bit for bit. */
int snl_scene_set_geometry(snl_scene *, const void *nodes, int nNodes, const void *tris, const void *shTris,
						   int nTris, const int *triMatId, int nMatIds);

/* Device-side replacement of BVH::Construct (src/bvh/tree.cpp:289-328) for a triangle soup: tris / shTris (may be
 * NULL) / triMatId as in snl_scene_set_geometry, in any order. Builds a Morton-order linear BVH on the GPU (radix
 * sort + Karras hierarchy + bottom-up refit with SAH leaf collapse, snail_b200/csrc/snl_build.cu) straight into the
 * traversal layout. NOT the reference's tree: hits are identical wherever they are unambiguous, triangle ids refer
 * to the builder's own order (use t / barycentrics / frames to compare). snl_scene_build_ms = device time of it. */
int snl_scene_build_device(snl_scene *, const void *tris, const void *shTris, int nTris, const int *triMatId,
						   int nMatIds);
float snl_scene_build_ms(snl_scene *);
/* Which device builder snl_scene_build_device (and SNL_BUILD_DEVICE) runs:
 *   SNL_BUILDER_LBVH  Morton-order linear BVH: ~0.4 ms for 268 K triangles, frames 13-25 % slower than on the SAH tree;
 *   SNL_BUILDER_SAH   top-down 16-bin surface-area-heuristic build - the algorithm of the reference's BVH::FindSplit
 *                     (src/bvh/tree.cpp:161-287: 16 bins per axis, costs traverse 0 / intersect 1, leaves of <= 4
 *                     triangles) as a level-synchronous GPU build (snail_b200/csrc/snl_build_sah.cu): every level is a few
 *                     flat kernels over triangles / nodes with integer min / max / add atomics (~4.6 ms for 268 K
 *                     triangles, frames within 2.3 % of the host SAH tree's), so the tree does not
 *                     depend on scheduling and the same code runs as plain loops on the host (snl_host_build_binned_sah). */
enum { SNL_BUILDER_LBVH = 0, SNL_BUILDER_SAH = 1 };
int snl_scene_set_device_builder(snl_scene *, int builder);
/* The binned-SAH build on the HOST (no device): tris = Triangle[nTris] (reference layout); gpuNodes receives up to
 * max(nTris - 1, 1) 64-byte traversal nodes {child0 minx,maxx,miny,maxy | child1 ... | c0 minz,maxz,c1 minz,maxz | link0,
 * link1, 0, 0} (link >= 0: node index, < 0: ~(first | (count - 1) << 26) into the re-ordered triangles), order[i] = index
 * of the input triangle that becomes triangle i. */
int snl_host_build_binned_sah(const void *tris, int nTris, void *gpuNodes, int *nNodes, int *order, int *height);

/* Instanced two-level BVH: Scene<DBVH> over ObjectInstances of this scene's BVH (src/dbvh/tree.h, the `-instances n`
 * mode of src/rtracer.cpp:146-178,359-363). snl_instance = the data members of ObjectInstance (:177-178): rotation rows
 * (orthonormal), translation. Computes every instance's box (ObjectInstance::ComputeBBox), builds the top-level tree
 * with DBVH::Construct's binned SAH (src/dbvh/tree.cpp:40-173) - same nodes, same element order as the reference - and
 * from then on snl_render* and snl_trace_primary_inst walk top tree -> instance -> BVH on the device. Like the
 * reference's Scene<DBVH> (HasShadingData() == 0) an instanced scene is shaded with the simple path: geometric normal
 * rotated into world space, default material; shadows and the reflection bounce work. n = 0 returns to the plain BVH.
 * snl_trace_rays / snl_trace_shadow / snl_raytrace and gVals[9] return SNL_ERR_UNSUPPORTED while instanced. New geometry
 * (snl_scene_set_geometry / snl_scene_build_device / snl_scene_load_obj) drops the instances: set them again after it. */
typedef struct snl_instance {
	float rotation[9]; /* rotation[0..2].xyz */
	float translation[3];
} snl_instance;
int snl_scene_set_instances(snl_scene *, const snl_instance *instances, int n);
/* DBVH::TraversePrimary(Context<1,0>&): instId = Context::object (index into DBVH::elements, i.e. the builder's
 * order, see snl_scene_copy_dbvh), triId = Context::element. */
int snl_trace_primary_inst(snl_scene *, const snl_camera *cam, int W, int H, int *instId, int *triId, float *t,
						   float *bary2);
/* DBVH::nodes (32-byte DBVH::Node records) and DBVH::elements in the builder's order, with their boxes {min, max}.
 * Any out pointer may be NULL. */
int snl_scene_copy_dbvh(snl_scene *, int *nNodes, void *nodes, int *nInstances, snl_instance *elements, float *bboxes6);
/* The same host builder without a device (ComputeBBox + DBVH::Construct over n instances of a tree with the given
 * box): nodesOut holds up to 2n DBVH::Node records, elementsOut / bboxes6Out n entries in the builder's order. */
int snl_host_dbvh_build(const snl_instance *instances, int n, const float treeMin[3], const float treeMax[3],
						void *nodesOut, int *nNodes, snl_instance *elementsOut, float *bboxes6Out);

/* Scene::materials / texDict (src/scene.h:25-40, src/scene.cpp:12-21). */
int snl_scene_set_materials(snl_scene *, const snl_material *mats, int nMats, const snl_texture *tex, int nTex);

/* Scene::lights + Scene::ambientLight (src/scene.h:49-50, src/scene.cpp:9). ambient may be NULL (= 0.1,0.1,0.1). */
int snl_scene_set_lights(snl_scene *, const snl_light *lights, int nLights, const float ambient[3]);

/* The global `int gVals[16]` (src/rtbase.cpp:16). Implemented: [1] depth shading, [6] full shading,
 * [7] reflections, [8] rank tint (uses `rank` of snl_render), [9] 2x2 supersampling in snl_render*
 * (src/render.cpp:71-110). [5] (stats heat-map) returns SNL_ERR_UNSUPPORTED. */
int snl_set_options(snl_scene *, const int gVals[16]);

/* Reflection bounces followed when gVals[7] is on. 1 (default) = the reference: `cache.reflections < 1`
 * (src/scene_trace.cpp:454-460) lets the bounce pass shade and light its hits but not bounce again. 2 = that condition
 * relaxed to `< 2` (BASELINE config 5, "reflection depth 2"): BEYOND the reference, compared in the tests with the
 * reference copy oracle/build_ref.sh patches the same way (oracle/_ref/libsnailref_depth2.so). Plain scenes only
 * (instanced scenes, transparent materials and gVals[9] return SNL_ERR_UNSUPPORTED at render time); snl_raytrace keeps
 * depth 1. */
int snl_set_reflection_depth(snl_scene *, int depth);

/* Host ingest + SAH builder (host side of the path): BaseScene::LoadWavefrontObj
 * (src/formats/wavefront_obj.cpp:66-200), FlipNormals/GenNormals (src/base_scene.cpp:326-335,517-529),
 * BVH::Construct (src/bvh/tree.cpp:293-328), LoadMaterialDescs/LoadTextures/MakeMaterials
 * (src/shading/material.cpp:59-192), exactly the set-up of src/rtracer.cpp:505-586; then the flatten step.
 * flipNormalsFlag / buildFlags as conf.flipNormals / conf.buildFlags there (flag 0 => normals ARE flipped;
 * buildFlags: 1 noShadingData, 2 fastBuild, 4 useSah; ours: 8 = SNL_BUILD_DEVICE, the host only parses the file and
 * the tree is built on the GPU, see snl_scene_build_device). mtlPath/texDir may be NULL. */
#define SNL_BUILD_DEVICE 8
#define SNL_BUILD_SAH_BINNED 16 /* with SNL_BUILD_DEVICE: the binned-SAH device builder instead of the LBVH */
int snl_scene_load_obj(snl_scene *, const char *objPath, const char *mtlPath, const char *texDir,
					   int flipNormalsFlag, int buildFlags);

/* The same host ingest + builder without a device (host logic only; usable on a machine with no GPU), and the
 * hand-over of its result to a scene (= the flatten + upload step of snl_scene_load_obj). */
typedef struct snl_host_scene snl_host_scene;
snl_host_scene *snl_host_scene_load_obj(const char *objPath, const char *mtlPath, const char *texDir,
										int flipNormalsFlag, int buildFlags);
void snl_host_scene_free(snl_host_scene *);
int snl_host_scene_counts(const snl_host_scene *, int *nNodes, int *nTris, int *nMatIds, int *depth, int *nMaterials,
						  int *nTextures);
/* reference layouts: BVH::Node[nNodes], Triangle[nTris], ShTriangle[nTris] (untouched when built with
 * noShadingData), BVH::materials[i].id */
int snl_host_scene_copy(const snl_host_scene *, void *nodes, void *tris, void *shTris, int *triMatId);
int snl_host_scene_material(const snl_host_scene *, int idx, snl_material *out);
int snl_host_scene_texture(const snl_host_scene *, int idx, snl_texture *out); /* out->data stays owned by the handle */
int snl_scene_set_host_scene(snl_scene *, const snl_host_scene *);

/* Read back what the host builder produced (reference layouts, for parity checks against the reference's
 * BVH::nodes/tris/shTris). Any out pointer may be NULL. */
int snl_scene_counts(snl_scene *, int *nNodes, int *nTris, int *nMatIds, int *depth, int *nMaterials);
int snl_scene_copy_host_bvh(snl_scene *, void *nodes, void *tris, void *shTris, int *triMatId);
int snl_scene_bbox(snl_scene *, float bmin[3], float bmax[3]); /* BVH::GetBBox() = nodes[0].bbox */

/* BVH::TraversePrimary(Context<1,0>&) on the primary rays RayGenerator::Generate produces for the full WxH image
 * (src/ray_generator.cpp:23-47, src/render.cpp:61-62,112-115, src/bvh/traverse.cpp:14-80). Outputs (host, each
 * may be NULL) are row-major per pixel: triId (-1 = miss), t (+inf = miss), bary2 = Context::barycentric (u,v). */
int snl_trace_primary(snl_scene *, const snl_camera *cam, int W, int H, int *triId, float *t, float *bary2);

/* BVH::TraversePrimary(Context<sharedOrigin,hasMask>&) on caller rays (AoS xyz). o3 holds one origin when
 * sharedOrigin, else n. mask (n bytes, 0 = inactive) may be NULL. Inactive / missed rays: triId -1, t +inf. */
int snl_trace_rays(snl_scene *, const float *o3, const float *d3, const unsigned char *mask, int n,
				   int sharedOrigin, int *triId, float *t, float *bary2);

/* BVH::TraverseShadow(ShadowContext&) (src/bvh/traverse.cpp:82-149): rays from one origin; tmax_inout[i] < 0 on
 * input = inactive; on output occluded rays hold -inf, the others keep their value. */
int snl_trace_shadow(snl_scene *, const float origin[3], const float *d3, float *tmax_inout, int n);

/* Scene<BVH>::RayTrace(const RayGroup<sharedOrigin,hasMask>&, Cache&, Vec3q *outColor)
 * (src/scene.h:44-46, src/scene_trace.cpp:92-520) on consecutive 256-ray packets of caller rays: ray i of a
 * packet is lane i&3 of quad i>>2. The last packet is padded with masked-off rays. rgb3 = outColor.{x,y,z}. */
int snl_raytrace(snl_scene *, const float *o3, const float *d3, const unsigned char *mask, int n,
				 int sharedOrigin, float *rgb3, snl_stats *stats);

/* Render() (src/render.h:16-23, src/render.cpp:214-280): the whole frame, or the tile list of the farm overload
 * (tiles_xywh = coords, 4 ints per tile; NULL = one tile covering the image). Pixels are written at
 * out + y*pitch + 3*x (absolute image coordinates, pitch in bytes) in ConvColor's byte order
 * (src/render.cpp:11-17: byte0 = colour.z, byte1 = colour.y, byte2 = colour.x); pixels outside the tiles are
 * left untouched. Tile x and w must be multiples of 16 (the packet width, src/render.cpp:52-53).
 * snl_render: host output buffer (device->host copy included). snl_render_dev: device output buffer, asynchronous
 * on the scene's stream until snl_scene_sync. rank only feeds the gVals[8] tint. */
int snl_render(snl_scene *, const snl_camera *cam, int W, int H, const int *tiles_xywh, int nTiles,
			   unsigned char *bgr8_host, size_t pitch, int rank, snl_stats *stats);
int snl_render_dev(snl_scene *, const snl_camera *cam, int W, int H, const int *tiles_xywh, int nTiles,
				   void *bgr8_dev, size_t pitch, int rank);
int snl_scene_sync(snl_scene *, snl_stats *stats);

/* The farm overload of Render() with its wire format (src/render.h:20-23, src/render.cpp:243-265, RenderTask with
 * compress == 1, src/render.cpp:140-166): tile i = coords[4i..4i+3] = {x, y, w, h} is written at data + offsets[i] as
 * three w*h byte planes - colour.x, colour.y - colour.x, colour.z - colour.x (mod 256, what DecompressTask of
 * src/compression.cpp:132-139 undoes) - w bytes per row. The reference only writes tiles of exactly 16 x 64 (its
 * maxWidth x maxHeight; others hit a TODO and stay unwritten); here any tile with x, w multiples of 16 is written in the
 * same layout. data is a host buffer of at least max(offsets[i] + 3*w*h) bytes; bytes between tiles are untouched.
 * rank feeds the gVals[8] tint like the reference's. The header / QuickLZ framing of src/node.cpp is not part of this. */
int snl_render_farm(snl_scene *, const snl_camera *cam, int resx, int resy, unsigned char *data, const int *coords,
					const int *offsets, int nTiles, int rank, snl_stats *stats);

/* Multi-GPU compose over NVLink peer memory (replaces the tile transport of src/node.cpp:20-120 / src/comm_mpi.cpp and
 * the per-tile receive loop of src/server.cpp:330-418). Rank 0 creates the frame on its device and hands the 64-byte
 * handle to the other ranks (any byte transport: torch.distributed broadcast in bench.py); each of them maps the SAME
 * HBM buffer with snl_frame_open and passes the mapped pointer to snl_render_dev together with its own tile list:
 * k_render then stores every pixel directly into rank 0's frame through NVSwitch - tracing and gather are one kernel,
 * no staging copy and no collective on the data path. After every rank's snl_scene_sync (+ a host barrier) the frame is
 * complete; snl_frame_read copies it to the host. pitch of a frame is W * 3. */
#define SNL_IPC_HANDLE_BYTES 64
int snl_frame_create(snl_scene *, int W, int H, void **bgr8_dev, unsigned char ipc_handle[SNL_IPC_HANDLE_BYTES]);
int snl_frame_open(snl_scene *, const unsigned char ipc_handle[SNL_IPC_HANDLE_BYTES], void **bgr8_dev);
int snl_frame_close(snl_scene *, void *bgr8_dev); /* frees (owner) or unmaps (peer) */
int snl_frame_read(snl_scene *, const void *bgr8_dev, int W, int H, unsigned char *bgr8_host, size_t host_pitch);

/* Render() over the GPUs of one box behind this ABI: replaces the reference's farm loop (src/server.cpp:227-265 tile
 * split, :330-418 receive loop + stats merge; src/node.cpp:308-346 node side; src/comm_mpi.cpp transport). Every GPU
 * holds a replica of the scene (one snl_scene each, same geometry / materials / lights / options); the frame lives in
 * rank 0's HBM in a segment the other ranks map. The frame's packets are dealt to the ranks like the reference's taskMap
 * (tiles of 4 packets, every run of `world` tiles in a freshly shuffled order; snl_comm_plan returns the split). Per
 * frame every rank runs one k_render per row band of its share, storing its packets straight into rank 0's frame over
 * NVLink, and signals the band with ONE system-scope increment behind the kernel boundary; rank 0 copies band i to the
 * host while band i+1 is still rendering on all GPUs, and sums the TreeStats. From the second frame on a band's packets
 * run longest first (cycles measured in the previous frame). No host-side collective per frame; every device-side wait
 * has an 8 s time-out that comes back as SNL_ERR_STATE instead of a hang (the communicator is then dead: destroy it).
 * A cursor-fed persistent k_render was built and measured slower.
 *   between processes (one rank per process, e.g. under torchrun / mpirun):
 *       rank 0:  snl_comm_create_root(scene, world, W, H, blob, &comm);  send blob to the others with ANY byte transport
 *       rank r:  snl_comm_attach(scene, r, blob, &comm);
 *   inside one process driving n devices:  snl_comm_create_local(scenes, n, W, H, &comm)   (scenes[0] = rank 0)
 * snl_render_multi is collective: every rank calls it once per frame with the same camera. bgr8_host (rank 0 only,
 * may be NULL to keep the frame on the device, see snl_comm_frame) receives the frame as snl_render writes it. stats:
 * rank 0 gets the sum over the ranks, another process its own share. gVals[9] is not available here. */
typedef struct snl_comm snl_comm;
#define SNL_COMM_BLOB_BYTES 96
int snl_comm_create_root(snl_scene *, int world, int W, int H, unsigned char blob[SNL_COMM_BLOB_BYTES], snl_comm **out);
int snl_comm_attach(snl_scene *, int rank, const unsigned char blob[SNL_COMM_BLOB_BYTES], snl_comm **out);
int snl_comm_create_local(snl_scene *const *scenes, int n, int W, int H, snl_comm **out);
int snl_comm_info(const snl_comm *, int *world, int *W, int *H, int *localRanks);
/* The work split of snl_render_multi without a device (host logic only): the packets {x, y, w, h} rank `rank` of `world`
 * renders of a W x H frame, row bands in order; bandFirst[b] .. bandFirst[b + 1] (9 entries) = its packets of band b.
 * packets_xywh may be NULL to ask for the count (at most ceil(W/16) * ceil(H/16) packets). */
int snl_comm_plan(int world, int W, int H, int rank, int *packets_xywh, int *nPackets, int *nBands, int *bandFirst);
int snl_render_multi(snl_comm *, const snl_camera *cam, unsigned char *bgr8_host, size_t pitch, snl_stats *stats);
/* device pointer (rank 0's address space) + pitch of the frame the last snl_render_multi produced */
int snl_comm_frame(snl_comm *, void **bgr8_dev, size_t *pitch);
void snl_comm_destroy(snl_comm *);

/* Builder-side ceilings for the roofline of the traversal kernels (bench.py `roofline`): the reference's scenes are
 * L2-resident on a B200 and the walk is served by L1, so what bounds k_render is instruction issue and L1 / L2
 * bandwidth, not HBM. Streaming microbenchmarks timed with CUDA events on `device` (snail_b200/csrc/snl_peaks.cu);
 * allocates 2 GiB for a second. Replaces nothing in the reference (its analogue is the `it:` / `in:` statistics of
 * src/tree_stats.cpp:35-36). */
typedef struct snl_peaks {
	float l1_gbs;                 /* L1 -> SM, coalesced 128-bit loads of an L1-resident window, all SMs */
	float l2_gbs;                 /* L2 -> SM, 48 MB streamed with L1 bypassed */
	float hbm_read_gbs;           /* HBM -> SM, 2 GiB read once */
	float issue_gwarpinst_s;      /* measured: FFMA + integer ALU interleave, 1e9 warp instructions / s */
	float issue_peak_gwarpinst_s; /* SMs x 4 schedulers x max SM clock */
	float sm_clock_mhz_max;
	int sms;
} snl_peaks;
int snl_measure_peaks(int device, snl_peaks *out);

/* Number of kernels this library launched on the scene since creation (bench.py's gpu_launches). */
uint64_t snl_scene_launch_count(snl_scene *);

/* Which register budget of k_render the scene settled on for its current workload (0: 64 registers / 4 CTAs per SM,
 * 1: 48 registers / 5 CTAs per SM; -1: not decided yet - the first full frame of a workload is rendered three times with
 * each and the faster is kept; SNL_RENDER_VARIANT=0|1 pins it). */
int snl_scene_render_variant(snl_scene *);

/* cudaStream_t of the scene, as an integer, so a caller can order its own work / CUDA events after ours. */
uint64_t snl_scene_stream(snl_scene *);

#ifdef __cplusplus
}
#endif
#endif
