// TEST INFRASTRUCTURE — link-time shim for oracle/_ref (see oracle/build_ref.sh).
//
// The reference's hot path links against ~25 non-GL libfwk translation units. Three symbols it
// needs live in libfwk files that pull in OpenGL / SDL / libdwarf, which this image does not have
// (and which this task must not probe for). They are restated here:
//   fwk::imageSize / imageRowSize   libfwk/src/gfx/gl_format.cpp:150-158 (only rgb8 / rgba8 are used
//                                   by MipmapTexture, src/mipmap_texture.h:34-47)
//   fwk::Backtrace                  libfwk/src/sys/backtrace.cpp (diagnostics only) -> empty
//   loadFont                        src/funcs.cpp:120-124 is not compiled; src/funcs.cpp's
//                                   BoxPointDistanceSq (src/funcs.cpp:8-55) is restated below because
//                                   funcs.cpp itself includes GL headers.

#include "rtbase.h"
#include <fwk/gfx/gl_format.h>
#include <fwk/sys/backtrace.h>

namespace fwk {

int imageSize(GlFormat format, int width, int height) {
	int bpp = format == GlFormat::rgb8 ? 3 : 4;
	return width * height * bpp;
}

int imageRowSize(GlFormat format, int width) {
	int bpp = format == GlFormat::rgb8 ? 3 : 4;
	return width * bpp;
}

// Diagnostics only: an empty backtrace (libfwk/src/sys/backtrace.cpp needs libdwarf/addr2line).
Backtrace Backtrace::get(int, void *, bool) { return Backtrace(); }
string Backtrace::format(Maybe<int>) const { return string(); }
void Backtrace::operator>>(TextFormatter &) const {}

} // namespace fwk

// src/funcs.cpp:57-118; only reachable from BaseScene::Object::FindOptimalTrans, which the hot path
// never calls (src/rtracer.cpp:559 has Optimize() commented out).
Matrix<Vec4f> Inverse(const Matrix<Vec4f> &) { FATAL("Inverse(): not part of the oracle build"); }

// Squared distance from a point to an AABB (0 inside); used for the per-packet light cull in
// Scene::RayTrace (src/scene_trace.cpp:495-499).
float BoxPointDistanceSq(const BBox &box, const Vec3f &point) {
	float sum = 0.0f;
	const float p[3] = {point.x, point.y, point.z};
	const float lo[3] = {box.min.x, box.min.y, box.min.z};
	const float hi[3] = {box.max.x, box.max.y, box.max.z};
	for(int a = 0; a < 3; a++) {
		if(p[a] < lo[a]) {
			float d = p[a] - lo[a];
			sum += d * d;
		}
		else if(p[a] > hi[a]) {
			float d = p[a] - hi[a];
			sum += d * d;
		}
	}
	return sum;
}
