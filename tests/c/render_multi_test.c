/* Plain C caller of libsnail_b200.so (no Python, no torch): renders frames on n GPUs of this box through
 * snl_comm_create_local + snl_render_multi and checks them byte for byte, and ray count for ray count, against the
 * single-GPU snl_render of the same cameras. What a C/C++ host like the reference's server loop
 * (src/server.cpp:330-418) would link against.
 *
 *   render_multi_test scene.obj [nGpus=all] [W=640 H=360] [frames=4]
 * exit status 0 = identical. */
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <string.h>

#include "snail_b200.h"

#define CHECK(call)                                                              \
	do {                                                                         \
		int r_ = (call);                                                         \
		if(r_ != SNL_OK) {                                                       \
			fprintf(stderr, "%s -> %d: %s\n", #call, r_, snl_last_error());     \
			return 2;                                                            \
		}                                                                        \
	} while(0)

int main(int argc, char **argv) {
	if(argc < 2) {
		fprintf(stderr, "usage: %s scene.obj [nGpus] [W H] [frames]\n", argv[0]);
		return 2;
	}
	int n = argc > 2 ? atoi(argv[2]) : snl_device_count();
	const int W = argc > 4 ? atoi(argv[3]) : 640, H = argc > 4 ? atoi(argv[4]) : 360;
	const int frames = argc > 5 ? atoi(argv[5]) : 4;
	if(n < 1 || n > snl_device_count()) {
		fprintf(stderr, "%d GPUs asked for, %d visible\n", n, snl_device_count());
		return 2;
	}
	snl_scene *scenes[16];
	int gv[16] = {0};
	gv[6] = 1;
	gv[7] = 1;
	float bmin[3], bmax[3];
	for(int i = 0; i < n; i++) {
		scenes[i] = snl_scene_create(i);
		if(!scenes[i]) {
			fprintf(stderr, "snl_scene_create(%d): %s\n", i, snl_last_error());
			return 2;
		}
		CHECK(snl_scene_load_obj(scenes[i], argv[1], NULL, NULL, 0, 4));
		CHECK(snl_set_options(scenes[i], gv));
	}
	CHECK(snl_scene_bbox(scenes[0], bmin, bmax));
	float c[3], size = 0.0f;
	for(int a = 0; a < 3; a++) {
		c[a] = 0.5f * (bmin[a] + bmax[a]);
		size += (bmax[a] - bmin[a]) * (bmax[a] - bmin[a]);
	}
	size = sqrtf(size);
	snl_comm *comm = NULL;
	CHECK(snl_comm_create_local(scenes, n, W, H, &comm));
	unsigned char *single = malloc((size_t)W * H * 3), *multi = malloc((size_t)W * H * 3);
	int bad = 0;
	for(int f = 0; f < frames; f++) {
		/* a camera that moves every frame: per-origin records are rebuilt, both frame slots get used */
		snl_camera cam;
		memset(&cam, 0, sizeof(cam));
		cam.pos[0] = c[0] + 0.05f * size * (float)(f % 3);
		cam.pos[1] = c[1] + 0.02f * size * (float)f;
		cam.pos[2] = c[2] - 0.75f * size;
		cam.right[0] = cam.up[1] = cam.front[2] = 1.0f;
		cam.plane_dist = 1.0f;
		snl_light light;
		memcpy(light.pos, cam.pos, 12);
		light.color[0] = light.color[1] = light.color[2] = 1.0f;
		light.radius = size;
		light.radSq = size * size;
		light.iRadius = 1.0f / size;
		for(int i = 0; i < n; i++)
			CHECK(snl_scene_set_lights(scenes[i], &light, 1, NULL));
		snl_stats s1, sn;
		memset(single, 0x55, (size_t)W * H * 3);
		memset(multi, 0xaa, (size_t)W * H * 3);
		CHECK(snl_render(scenes[0], &cam, W, H, NULL, 0, single, (size_t)W * 3, 0, &s1));
		CHECK(snl_render_multi(comm, &cam, multi, (size_t)W * 3, &sn));
		const int same = memcmp(single, multi, (size_t)W * H * 3) == 0;
		printf("frame %d: %d GPU(s) %s the single-GPU frame; rays %llu vs %llu; %.3f ms vs %.3f ms\n", f, n,
			   same ? "byte-identical to" : "DIFFERS from", (unsigned long long)sn.rays, (unsigned long long)s1.rays, sn.kernel_ms,
			   s1.kernel_ms);
		bad += !same || sn.rays != s1.rays || sn.primary != s1.primary || sn.reflection_shadow != s1.reflection_shadow;
	}
	snl_comm_destroy(comm);
	for(int i = 0; i < n; i++)
		snl_scene_destroy(scenes[i]);
	free(single);
	free(multi);
	printf("%s\n", bad ? "FAIL" : "PASS");
	return bad ? 1 : 0;
}
