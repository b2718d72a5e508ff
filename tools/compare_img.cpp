// compare_img — the image comparator the parity criterion is stated in (north_star: "final images must pass
// tools/compare_img at max channel error <= 1/255 on >= 99.9 % of pixels").
//
//   compare_img image1 image2 [--tol N] [--frac F] [--quiet]
//
// Same CLI as the reference's tool (/root/reference/tools/compare_img.cpp:1-33: `compare_img image1 image2`, which
// only prints the mean R, G, B of each file); this one prints those two lines too and then DECIDES: exit status 0 iff
// the images have the same size and at least F (default 0.999) of the pixels have a max channel error <= N (default 1).
// Exit 1 = images differ, 2 = usage / unreadable file.
//
// Formats (by content, not by extension): binary PPM (P6, maxval 255), uncompressed true-colour TGA (24 / 32 bit, either
// origin), 24 / 32-bit uncompressed BMP. Channels are compared position by position; the tool does not care whether a
// file holds RGB or ConvColor's BGR order (src/render.cpp:11-17) as long as both files use the same.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct Img {
	int w = 0, h = 0;
	std::vector<unsigned char> px;  // 3 bytes per pixel, rows top to bottom, channel order as stored in the file
};

static bool read_file(const char *path, std::vector<unsigned char> &out) {
	FILE *f = fopen(path, "rb");
	if(!f)
		return false;
	fseek(f, 0, SEEK_END);
	long n = ftell(f);
	fseek(f, 0, SEEK_SET);
	out.resize(n > 0 ? (size_t)n : 0);
	bool ok = n >= 0 && fread(out.data(), 1, out.size(), f) == out.size();
	fclose(f);
	return ok;
}

static bool parse_ppm(const std::vector<unsigned char> &d, Img &im) {
	if(d.size() < 2 || d[0] != 'P' || d[1] != '6')
		return false;
	size_t p = 2;
	int vals[3], got = 0;
	while(got < 3 && p < d.size()) {
		if(d[p] == '#') {
			while(p < d.size() && d[p] != '\n')
				p++;
		}
		else if(d[p] == ' ' || d[p] == '\t' || d[p] == '\n' || d[p] == '\r')
			p++;
		else if(d[p] >= '0' && d[p] <= '9') {
			long v = 0;
			while(p < d.size() && d[p] >= '0' && d[p] <= '9' && v < (1 << 28))
				v = v * 10 + (d[p++] - '0');
			vals[got++] = (int)v;
		}
		else
			return false;
	}
	if(got < 3 || vals[2] != 255 || p >= d.size())
		return false;
	p++;  // the single whitespace byte after maxval
	im.w = vals[0];
	im.h = vals[1];
	const size_t need = (size_t)im.w * im.h * 3;
	if(im.w <= 0 || im.h <= 0 || d.size() - p < need)
		return false;
	im.px.assign(d.begin() + p, d.begin() + p + need);
	return true;
}

static bool parse_tga(const std::vector<unsigned char> &d, Img &im) {
	if(d.size() < 18 || d[1] != 0 || d[2] != 2)
		return false;
	const int idLen = d[0], bpp = d[16];
	im.w = d[12] | d[13] << 8;
	im.h = d[14] | d[15] << 8;
	if((bpp != 24 && bpp != 32) || im.w <= 0 || im.h <= 0)
		return false;
	const int bytes = bpp / 8;
	const size_t off = 18 + (size_t)idLen;
	if(d.size() < off + (size_t)im.w * im.h * bytes)
		return false;
	const bool topDown = (d[17] & 0x20) != 0;
	im.px.resize((size_t)im.w * im.h * 3);
	for(int y = 0; y < im.h; y++) {
		const unsigned char *src = d.data() + off + (size_t)(topDown ? y : im.h - 1 - y) * im.w * bytes;
		unsigned char *dst = im.px.data() + (size_t)y * im.w * 3;
		for(int x = 0; x < im.w; x++)
			memcpy(dst + x * 3, src + x * bytes, 3);
	}
	return true;
}

static bool parse_bmp(const std::vector<unsigned char> &d, Img &im) {
	if(d.size() < 54 || d[0] != 'B' || d[1] != 'M')
		return false;
	auto u32 = [&](size_t o) { return (unsigned)d[o] | (unsigned)d[o + 1] << 8 | (unsigned)d[o + 2] << 16 | (unsigned)d[o + 3] << 24; };
	const unsigned off = u32(10);
	const int w = (int)u32(18), hs = (int)u32(22);
	const int bpp = d[28] | d[29] << 8;
	if(u32(30) != 0 || (bpp != 24 && bpp != 32) || w <= 0 || hs == 0)
		return false;
	im.w = w;
	im.h = hs < 0 ? -hs : hs;
	const int bytes = bpp / 8;
	const size_t row = ((size_t)w * bytes + 3) & ~(size_t)3;
	if(d.size() < off + row * im.h)
		return false;
	im.px.resize((size_t)im.w * im.h * 3);
	for(int y = 0; y < im.h; y++) {
		const unsigned char *src = d.data() + off + (size_t)(hs < 0 ? y : im.h - 1 - y) * row;
		unsigned char *dst = im.px.data() + (size_t)y * im.w * 3;
		for(int x = 0; x < im.w; x++)
			memcpy(dst + x * 3, src + x * bytes, 3);
	}
	return true;
}

static bool load(const char *path, Img &im) {
	std::vector<unsigned char> d;
	if(!read_file(path, d)) {
		fprintf(stderr, "compare_img: cannot read %s\n", path);
		return false;
	}
	if(parse_ppm(d, im) || parse_bmp(d, im) || parse_tga(d, im))
		return true;
	fprintf(stderr, "compare_img: %s is not a P6 PPM / uncompressed TGA / uncompressed BMP\n", path);
	return false;
}

int main(int argc, char **argv) {
	int tol = 1;
	double frac = 0.999;
	bool quiet = false;
	std::vector<const char *> files;
	for(int i = 1; i < argc; i++) {
		if(!strcmp(argv[i], "--tol") && i + 1 < argc)
			tol = atoi(argv[++i]);
		else if(!strcmp(argv[i], "--frac") && i + 1 < argc)
			frac = atof(argv[++i]);
		else if(!strcmp(argv[i], "--quiet"))
			quiet = true;
		else
			files.push_back(argv[i]);
	}
	if(files.size() != 2) {
		printf("Usage\ncompare_img image1 image2 [--tol N=1] [--frac F=0.999] [--quiet]\n");
		return 2;
	}
	Img img[2];
	for(int n = 0; n < 2; n++)
		if(!load(files[n], img[n]))
			return 2;
	if(!quiet)
		for(int n = 0; n < 2; n++) {  // the reference tool's output: mean of every channel
			long long sum[3] = {0, 0, 0};
			for(size_t i = 0; i < img[n].px.size(); i += 3)
				for(int c = 0; c < 3; c++)
					sum[c] += img[n].px[i + c];
			const double icount = 1.0 / ((double)img[n].w * (double)img[n].h);
			printf("%s:\t\tR:%.2f\tG:%.2f\tB:%.2f\n", files[n], sum[0] * icount, sum[1] * icount, sum[2] * icount);
		}
	if(img[0].w != img[1].w || img[0].h != img[1].h) {
		printf("size mismatch: %dx%d vs %dx%d\nFAIL\n", img[0].w, img[0].h, img[1].w, img[1].h);
		return 1;
	}
	const size_t pixels = (size_t)img[0].w * img[0].h;
	size_t within = 0, identical = 0;
	int maxErr = 0;
	long long hist[4] = {0, 0, 0, 0};  // per-pixel max channel error 0, 1, 2..8, > 8
	for(size_t p = 0; p < pixels; p++) {
		int e = 0;
		for(int c = 0; c < 3; c++) {
			const int d = abs((int)img[0].px[p * 3 + c] - (int)img[1].px[p * 3 + c]);
			e = d > e ? d : e;
		}
		within += e <= tol;
		identical += e == 0;
		maxErr = e > maxErr ? e : maxErr;
		hist[e == 0 ? 0 : e == 1 ? 1 : e <= 8 ? 2 : 3]++;
	}
	const double f = (double)within / (double)pixels;
	const bool pass = f >= frac;
	if(!quiet)
		printf("%dx%d: max channel error <= %d on %.5f %% of pixels (need %.3f %%); identical %.5f %%; max error %d; "
			   "histogram 0:%lld 1:%lld 2-8:%lld >8:%lld\n",
			   img[0].w, img[0].h, tol, f * 100.0, frac * 100.0, (double)identical / pixels * 100.0, maxErr, hist[0], hist[1],
			   hist[2], hist[3]);
	printf("%s\n", pass ? "PASS" : "FAIL");
	return pass ? 0 : 1;
}
