// TEST INFRASTRUCTURE — not part of the product.
//
// Compiles the product's reference-side adapter (include/scene_b200.h: `Scene<B200Acc>`, RayTrace on the B200) INSIDE
// the reference, against the reference's own headers, and instantiates the reference's Render() templates for it:
// src/render.cpp is included where it lies (its templates are defined in the .cpp), so the frame below is produced by
// the reference's own RenderTask::Work packet loop (src/render.cpp:47-211), its ray generator, thread pool and
// ConvColor, with ONLY `scene->RayTrace(...)` (src/render.cpp:81,115) answered by libsnail_b200.so.
// Built by oracle/build_ref.sh into its own library, oracle/_ref/libsnailref_b200adapter.so (the plain oracle libraries
// never link the product), and driven by tests/test_gpu_adapter.py.
#include "scene_b200.h"

#include "render.cpp"  // resolved through -I$REF/src

template TreeStats Render<B200Acc>(const Scene<B200Acc> &, const Camera &, MipmapTexture &, const Options, uint);

#include <chrono>

namespace {
struct RefSceneView {  // the leading members of oracle/ref_capi.cpp's RefScene
	Scene<BVH> scene;
	vector<shading::MaterialDesc> matDescs;
};
struct Adapter {
	SceneB200 scene;
	Adapter(const RefSceneView &rs, int device) : scene(rs.scene, rs.matDescs, device) {}
};
}  // namespace

extern "C" {

// h = a handle of ref_scene_load (oracle/ref_capi.cpp) of THIS library. Returns 0 on failure (message on stderr).
void *ref_adapter_create(void *h, int device) {
	try {
		return new Adapter(*(const RefSceneView *)h, device);
	} catch(const std::exception &e) {
		fprintf(stderr, "ref_adapter_create: %s\n", e.what());
		return 0;
	}
}

void ref_adapter_free(void *a) { delete(Adapter *)a; }

// lights: n x {pos.xyz, color.rgb, radius}
void ref_adapter_set_lights(void *a, const float *l, int n) {
	auto &lights = ((Adapter *)a)->scene.lights;
	lights.clear();
	for(int i = 0; i < n; i++, l += 7)
		lights.push_back(Light(Vec3f(l[0], l[1], l[2]), Vec3f(l[3], l[4], l[5]), l[6]));
}

// The reference's Render(scene, camera, image, options, threads) with AccStruct = B200Acc. mode 1: the adapter's batch
// entry point (SceneB200::RenderFrame = one snl_render) instead. Returns wall milliseconds, < 0 on failure.
double ref_adapter_render(void *a, const float *c, int W, int H, int threads, int mode, unsigned char *out, unsigned *raysOut) {
	Adapter *ad = (Adapter *)a;
	Camera camera(Vec3f(c[0], c[1], c[2]), Vec3f(c[3], c[4], c[5]), Vec3f(c[6], c[7], c[8]), Vec3f(c[9], c[10], c[11]), c[12]);
	MipmapTexture image(W, H, fwk::GlFormat::rgb8, 1);
	memset(image.DataPointer(), 0, image.Size());
	auto t0 = std::chrono::steady_clock::now();
	TreeStats st;
	try {
		st = mode == 1 ? ad->scene.RenderFrame(camera, image) : Render(ad->scene, camera, image, Options(), (uint)threads);
	} catch(const std::exception &e) {
		fprintf(stderr, "ref_adapter_render: %s\n", e.what());
		return -1.0;
	}
	auto t1 = std::chrono::steady_clock::now();
	for(int y = 0; y < H; y++)
		memcpy(out + (size_t)y * W * 3, image.DataPointer() + image.Pitch() * y, (size_t)W * 3);
	if(raysOut)
		*raysOut = st.GetRays();
	return std::chrono::duration<double, std::milli>(t1 - t0).count();
}

}  // extern "C"
