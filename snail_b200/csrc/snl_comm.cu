// snail_b200 — Render() over the GPUs of one box behind the C ABI (snl_comm_*, snl_render_multi): the replacement of the
// reference's render farm loop (src/server.cpp:227-265 tile split, :330-418 per-frame receive loop and stats merge;
// src/node.cpp:308-346 the node side; src/comm_mpi.cpp transport).
//
// The scene is replicated (one snl_scene per GPU); the frame lives in rank 0's HBM inside ONE shared segment that the
// other ranks map (CUDA IPC between processes, peer access inside one process):
//
//     [ CommCtl (4 KB) | frame slot 0 | frame slot 1 ]
//
// Work split: the frame's packets are dealt to the ranks in tiles of 4 packets (64 x 16 pixels), every run of `world`
// consecutive tiles going to all ranks in a freshly shuffled order (the reference's taskMap, src/server.cpp:238-248),
// and each rank's list is kept in row-band order. Per frame `seq` (slot = seq & 1), everything in-stream, no host-side
// collective:
//   rank 0   k_frame_open     waits until every rank has left frame seq-2, zeroes the slot's counters, opens it
//   rank r   k_wait_counter   (r > 0) waits for the slot to be open
//   all      per row band, on the band's own stream:  k_render of the rank's packets of that band, storing 16-byte
//            vectors straight into rank 0's frame over NVLink, then k_band_signal: the kernel boundary has completed the
//            launch's stores, so ONE system-scope increment per (rank, band) says the share is in - no fence and no
//            atomic per packet (measured: a cursor-fed k_render with a release per packet cost 3 % on the headline and
//            15-20 % on the cheap 4K frames on one GPU, and ~10 us per packet on a peer)
//   rank 0   on the same band stream: k_wait_counter(band counter == world) + cudaMemcpy2DAsync of the band's rows: band
//            i goes to the host while band i+1 is still being rendered on every GPU
//   all      k_frame_publish  adds the rank's TreeStats counters to the frame's (src/server.cpp:403-418), marks the rank done
// Every wait has a time-out that raises CommCtl::error instead of hanging the device.
#include <cstdlib>
#include <cstring>

#include "snl_scene.h"

using namespace snl;

#define SNL_CUDA(call)                                                                              \
	do {                                                                                            \
		cudaError_t e_ = (call);                                                                    \
		if(e_ != cudaSuccess) {                                                                     \
			set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
			return SNL_ERR_CUDA;                                                                    \
		}                                                                                           \
	} while(0)
#define SNL_REQUIRE(cond, code, ...) \
	do {                             \
		if(!(cond)) {                \
			set_error(__VA_ARGS__);  \
			return code;             \
		}                            \
	} while(0)
#define SNL_TRY(expr)    \
	do {                 \
		int r_ = (expr); \
		if(r_)           \
			return r_;   \
	} while(0)

namespace {

struct CommRank {  // one rank driven by this process
	snl_scene *scene = nullptr;
	int rank = 0;
	unsigned char *seg = nullptr;  // rank 0's segment as this rank's device sees it
	bool ipc = false;              // mapped with cudaIpcOpenMemHandle (else: owned, or a peer pointer of this process)
	int4 *dPackets[2] = {nullptr, nullptr};  // this rank's packets {x, y, w, h}, row bands in order; [cur] is the live list
	int cur = 0;
	unsigned *dCost = nullptr;     // cycles every packet of the live list took in the last frame
	int *dBandFirst = nullptr;     // device copy of bandFirst
	int bandFirst[kCommBands + 1] = {};  // [b] .. [b + 1]: its packets of band b
	cudaEvent_t ev0 = nullptr, ev1 = nullptr;
	cudaStream_t sortStream = nullptr;  // the next frame's packet order is worked out beside the caller's stream
	cudaEvent_t sortIn = nullptr, sortOut = nullptr;
	bool sortPending = false;
};

struct Blob {  // what travels from rank 0 to the other processes (any byte transport)
	unsigned char handle[64];
	int world, W, H, magic;
};
static_assert(sizeof(Blob) <= SNL_COMM_BLOB_BYTES, "blob size");
enum { kBlobMagic = 0x534e4c43 };

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

struct snl_comm {
	int world = 0, W = 0, H = 0;
	size_t pitch = 0, frameBytes = 0, segBytes = 0;
	unsigned seq = 0;
	std::vector<CommRank> ranks;  // 1 entry between processes, `world` entries inside one process
	bool ownsSegment = false;
	// rank 0's side
	int nBands = 1, rowsPerBand = 1;
	bool sortByCost = true;  // next frame's packets of a band longest first (cycles measured this frame)
	CommCtl *hCtl = nullptr;  // pinned copy of the control block (summed counters, error word)
	float lastMs = 0.0f;
};

namespace {

int comm_geometry(snl_comm *c, int world, int W, int H) {
	SNL_REQUIRE(world >= 1 && world <= kCommRanks, SNL_ERR_INVALID, "world size %d (1..%d)", world, (int)kCommRanks);
	SNL_REQUIRE(W > 0 && H > 0, SNL_ERR_INVALID, "bad frame size %dx%d", W, H);
	c->world = world;
	c->W = W;
	c->H = H;
	c->pitch = align_up((size_t)W * 3, 16);
	c->frameBytes = align_up(c->pitch * H, 256);
	c->segBytes = kCommCtlBytes + 2 * c->frameBytes;
	const int px = (W + 15) / 16, py = (H + 15) / 16;
	// Row bands = launches per rank and frame (each band's rows go to the host while the next ones render). A band is
	// worth its launches (~10 us of host time) only while a rank has a good share of a wave of packets per band: 2 bands
	// for a 1080p frame on 8 GPUs (1020 packets per rank), 8 for a 4K frame.
	c->nBands = std::max(1, std::min<int>(std::min<int>(kCommBands, py), px * py / world / 500));
	if(const char *e = getenv("SNL_COMM_BANDS"))
		c->nBands = std::max(1, std::min<int>(std::min<int>(kCommBands, py), atoi(e)));
	// longest-packet-first within every band (measured on sponza_like 1080p, 2 GPUs: 0.780 -> 0.758 ms per frame); the
	// sorting kernel takes bands of up to 4096 packets
	c->sortByCost = px * py / world / c->nBands <= 3500;
	if(const char *e = getenv("SNL_COMM_SORT"))
		c->sortByCost = e[0] != '0';
	c->rowsPerBand = (py + c->nBands - 1) / c->nBands;
	c->nBands = (py + c->rowsPerBand - 1) / c->rowsPerBand;
	return SNL_OK;
}

int comm_root_resources(snl_comm *c) {
	CommRank &r0 = c->ranks[0];
	SNL_CUDA(cudaSetDevice(r0.scene->device));
	SNL_CUDA(cudaMallocHost(&c->hCtl, sizeof(CommCtl)));
	return SNL_OK;
}

int comm_rank_events(CommRank &r) {
	SNL_CUDA(cudaSetDevice(r.scene->device));

	SNL_CUDA(cudaEventCreate(&r.ev0));
	SNL_CUDA(cudaEventCreate(&r.ev1));
	SNL_CUDA(cudaStreamCreateWithFlags(&r.sortStream, cudaStreamNonBlocking));
	SNL_CUDA(cudaEventCreateWithFlags(&r.sortIn, cudaEventDisableTiming));
	SNL_CUDA(cudaEventCreateWithFlags(&r.sortOut, cudaEventDisableTiming));
	return ensure_band_streams(r.scene);
}

int alloc_segment(snl_comm *c, snl_scene *s, unsigned char **seg) {
	SNL_CUDA(cudaSetDevice(s->device));
	void *p = nullptr;
	SNL_CUDA(cudaMalloc(&p, c->segBytes));
	SNL_CUDA(cudaMemset(p, 0, c->segBytes));
	SNL_CUDA(cudaDeviceSynchronize());
	*seg = (unsigned char *)p;
	c->ownsSegment = true;
	return SNL_OK;
}

// The reference's taskMap (src/server.cpp:238-248): every run of `world` consecutive tiles goes to all ranks in a freshly
// shuffled order. A fixed generator, so every process computes the same map without talking.
std::vector<int> deal_tiles(int nTiles, int world) {
	std::vector<int> owner(nTiles), order(world);
	for(int i = 0; i < world; i++)
		order[i] = i;
	unsigned long long state = 0x9e3779b97f4a7c15ull;
	for(int n = 0; n < nTiles; n++) {
		if(n % world == 0)
			for(int i = world - 1; i > 0; i--) {  // Fisher-Yates
				state = state * 6364136223846793005ull + 1442695040888963407ull;
				std::swap(order[i], order[(int)((state >> 33) % (unsigned)(i + 1))]);
			}
		owner[n] = order[n % world];
	}
	return owner;
}

enum { kTilePackets = 4 };  // a tile = 4 packets of one packet row (64 x 16 pixels)

// The packets {x, y, w, h} of `rank`, row bands in order, and where each band starts in the list (host only).
void plan_packets(const snl_comm *c, int rank, std::vector<int4> &list, int bandFirst[kCommBands + 1]) {
	const int W = c->W, H = c->H, px = (W + 15) / 16, py = (H + 15) / 16;
	const int tilesX = (px + kTilePackets - 1) / kTilePackets;
	const std::vector<int> owner = deal_tiles(tilesX * py, c->world);
	list.clear();
	for(int b = 0; b < c->nBands; b++) {
		bandFirst[b] = (int)list.size();
		for(int row = b * c->rowsPerBand; row < std::min(py, (b + 1) * c->rowsPerBand); row++)
			for(int t = 0; t < tilesX; t++)
				if(owner[row * tilesX + t] == rank)
					for(int col = t * kTilePackets; col < std::min(px, (t + 1) * kTilePackets); col++)
						list.push_back(make_int4(col * 16, row * 16, std::min(16, W - col * 16), std::min(16, H - row * 16)));
	}
	for(int b = c->nBands; b <= kCommBands; b++)
		bandFirst[b] = (int)list.size();
}

int build_packet_list(snl_comm *c, CommRank &r) {
	std::vector<int4> list;
	plan_packets(c, r.rank, list, r.bandFirst);
	SNL_CUDA(cudaSetDevice(r.scene->device));
	const size_t n = std::max<size_t>(list.size(), 1);
	for(int k = 0; k < 2; k++)
		SNL_CUDA(cudaMalloc((void **)&r.dPackets[k], n * sizeof(int4)));
	SNL_CUDA(cudaMalloc((void **)&r.dCost, n * sizeof(unsigned)));
	SNL_CUDA(cudaMalloc((void **)&r.dBandFirst, sizeof(r.bandFirst)));
	SNL_CUDA(cudaMemset(r.dCost, 0, n * sizeof(unsigned)));
	if(!list.empty())
		SNL_CUDA(cudaMemcpy(r.dPackets[0], list.data(), list.size() * sizeof(int4), cudaMemcpyHostToDevice));
	SNL_CUDA(cudaMemcpy(r.dBandFirst, r.bandFirst, sizeof(r.bandFirst), cudaMemcpyHostToDevice));
	return SNL_OK;
}

}  // namespace

extern "C" {

int snl_comm_create_root(snl_scene *s, int world, int W, int H, unsigned char blob[SNL_COMM_BLOB_BYTES], snl_comm **out) {
	SNL_REQUIRE(s && out, SNL_ERR_INVALID, "null argument");
	snl_comm *c = new snl_comm;
	int r = comm_geometry(c, world, W, H);
	CommRank r0;
	r0.scene = s;
	r0.rank = 0;
	if(!r)
		r = alloc_segment(c, s, &r0.seg);
	c->ranks.push_back(r0);
	if(!r)
		r = comm_rank_events(c->ranks[0]);
	if(!r)
		r = build_packet_list(c, c->ranks[0]);
	if(!r)
		r = comm_root_resources(c);
	if(!r && blob) {
		Blob b;
		memset(&b, 0, sizeof(b));
		cudaIpcMemHandle_t h;
		cudaError_t e = cudaIpcGetMemHandle(&h, c->ranks[0].seg);
		if(e != cudaSuccess) {
			set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
			r = SNL_ERR_CUDA;
		}
		else {
			static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
			memcpy(b.handle, &h, 64);
			b.world = world;
			b.W = W;
			b.H = H;
			b.magic = kBlobMagic;
			memset(blob, 0, SNL_COMM_BLOB_BYTES);
			memcpy(blob, &b, sizeof(b));
		}
	}
	if(r) {
		snl_comm_destroy(c);
		return r;
	}
	*out = c;
	return SNL_OK;
}

int snl_comm_attach(snl_scene *s, int rank, const unsigned char blob[SNL_COMM_BLOB_BYTES], snl_comm **out) {
	SNL_REQUIRE(s && blob && out, SNL_ERR_INVALID, "null argument");
	Blob b;
	memcpy(&b, blob, sizeof(b));
	SNL_REQUIRE(b.magic == kBlobMagic, SNL_ERR_INVALID, "not a snl_comm blob");
	SNL_REQUIRE(rank > 0 && rank < b.world, SNL_ERR_INVALID, "rank %d outside 1..%d (rank 0 is the root)", rank, b.world - 1);
	snl_comm *c = new snl_comm;
	int r = comm_geometry(c, b.world, b.W, b.H);
	CommRank cr;
	cr.scene = s;
	cr.rank = rank;
	if(!r) {
		cudaIpcMemHandle_t h;
		memcpy(&h, b.handle, 64);
		void *p = nullptr;
		cudaError_t e = cudaSetDevice(s->device);
		if(e == cudaSuccess)
			e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
		if(e != cudaSuccess) {
			set_error("cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
			r = SNL_ERR_CUDA;
		}
		cr.seg = (unsigned char *)p;
		cr.ipc = p != nullptr;
	}
	c->ranks.push_back(cr);
	if(!r)
		r = comm_rank_events(c->ranks[0]);
	if(!r)
		r = build_packet_list(c, c->ranks[0]);
	if(r) {
		snl_comm_destroy(c);
		return r;
	}
	*out = c;
	return SNL_OK;
}

int snl_comm_create_local(snl_scene *const *scenes, int n, int W, int H, snl_comm **out) {
	SNL_REQUIRE(scenes && out && n >= 1, SNL_ERR_INVALID, "bad argument");
	for(int i = 0; i < n; i++) {
		SNL_REQUIRE(scenes[i], SNL_ERR_INVALID, "null scene %d", i);
		for(int j = 0; j < i; j++)
			SNL_REQUIRE(scenes[i]->device != scenes[j]->device, SNL_ERR_INVALID, "scenes %d and %d share device %d", j, i,
						scenes[i]->device);
	}
	snl_comm *c = new snl_comm;
	int r = comm_geometry(c, n, W, H);
	unsigned char *seg = nullptr;
	if(!r)
		r = alloc_segment(c, scenes[0], &seg);
	for(int i = 0; i < n && !r; i++) {
		CommRank cr;
		cr.scene = scenes[i];
		cr.rank = i;
		cr.seg = seg;  // unified addressing: a peer-enabled device dereferences rank 0's pointer directly
		c->ranks.push_back(cr);
		if(i > 0) {
			int can = 0;
			cudaDeviceCanAccessPeer(&can, scenes[i]->device, scenes[0]->device);
			if(!can) {
				set_error("device %d cannot access device %d's memory (no peer path)", scenes[i]->device, scenes[0]->device);
				r = SNL_ERR_UNSUPPORTED;
				break;
			}
			cudaSetDevice(scenes[i]->device);
			cudaError_t e = cudaDeviceEnablePeerAccess(scenes[0]->device, 0);
			if(e == cudaErrorPeerAccessAlreadyEnabled)
				cudaGetLastError();
			else if(e != cudaSuccess) {
				set_error("cudaDeviceEnablePeerAccess failed: %s", cudaGetErrorString(e));
				r = SNL_ERR_CUDA;
				break;
			}
		}
		r = comm_rank_events(c->ranks.back());
		if(!r)
			r = build_packet_list(c, c->ranks.back());
	}
	if(!r)
		r = comm_root_resources(c);
	if(r) {
		snl_comm_destroy(c);
		return r;
	}
	*out = c;
	return SNL_OK;
}

void snl_comm_destroy(snl_comm *c) {
	if(!c)
		return;
	for(CommRank &r : c->ranks) {
		if(!r.scene)
			continue;
		cudaSetDevice(r.scene->device);
		cudaStreamSynchronize(r.scene->stream);
		if(r.ev0)
			cudaEventDestroy(r.ev0);
		if(r.ev1)
			cudaEventDestroy(r.ev1);
		if(r.sortStream) {
			cudaStreamSynchronize(r.sortStream);
			cudaStreamDestroy(r.sortStream);
		}
		if(r.sortIn)
			cudaEventDestroy(r.sortIn);
		if(r.sortOut)
			cudaEventDestroy(r.sortOut);
		if(r.ipc && r.seg)
			cudaIpcCloseMemHandle(r.seg);
		for(int4 *p : r.dPackets)
			if(p)
				cudaFree(p);
		if(r.dCost)
			cudaFree(r.dCost);
		if(r.dBandFirst)
			cudaFree(r.dBandFirst);
	}
	if(!c->ranks.empty() && c->ranks[0].rank == 0 && c->ranks[0].scene) {
		cudaSetDevice(c->ranks[0].scene->device);
		if(c->hCtl)
			cudaFreeHost(c->hCtl);
		if(c->ownsSegment && c->ranks[0].seg)
			cudaFree(c->ranks[0].seg);
	}
	delete c;
}

int snl_comm_plan(int world, int W, int H, int rank, int *packets_xywh, int *nPackets, int *nBands, int *bandFirst) {
	SNL_REQUIRE(nPackets && rank >= 0 && rank < world, SNL_ERR_INVALID, "bad argument");
	snl_comm c;
	SNL_TRY(comm_geometry(&c, world, W, H));
	std::vector<int4> list;
	int first[kCommBands + 1];
	plan_packets(&c, rank, list, first);
	*nPackets = (int)list.size();
	if(nBands)
		*nBands = c.nBands;
	if(bandFirst)
		memcpy(bandFirst, first, sizeof(first));
	if(packets_xywh)
		memcpy(packets_xywh, list.data(), list.size() * sizeof(int4));
	return SNL_OK;
}

int snl_comm_info(const snl_comm *c, int *world, int *W, int *H, int *localRanks) {
	SNL_REQUIRE(c, SNL_ERR_INVALID, "null comm");
	if(world)
		*world = c->world;
	if(W)
		*W = c->W;
	if(H)
		*H = c->H;
	if(localRanks)
		*localRanks = (int)c->ranks.size();
	return SNL_OK;
}

int snl_comm_frame(snl_comm *c, void **bgr8_dev, size_t *pitch) {
	SNL_REQUIRE(c && bgr8_dev && c->seq > 0, SNL_ERR_STATE, "no frame rendered yet");
	*bgr8_dev = c->ranks[0].seg + kCommCtlBytes + (size_t)((c->seq - 1) & 1u) * c->frameBytes;
	if(pitch)
		*pitch = c->pitch;
	return SNL_OK;
}

int snl_render_multi(snl_comm *c, const snl_camera *cam, unsigned char *bgr8_host, size_t hostPitch, snl_stats *stats) {
	SNL_REQUIRE(c && cam, SNL_ERR_INVALID, "null argument");
	SNL_REQUIRE(!bgr8_host || hostPitch >= (size_t)c->W * 3, SNL_ERR_INVALID, "host pitch too small");
	const int W = c->W, H = c->H;
	const unsigned seq = c->seq, slot = seq & 1u;
	const int py = (H + 15) / 16;
	bool isRoot = false;
	// Phase 1, every local rank: whatever may allocate or synchronise (variant pick on the first frame, per-origin records).
	// With several devices in ONE process this must be over before any device starts waiting on a peer: a cudaMalloc on
	// one device updates its peers' mappings and would block behind their waiting kernels.
	std::vector<SceneDev> devScene(c->ranks.size());
	std::vector<RayGenParams> rayGen(c->ranks.size());
	for(size_t i = 0; i < c->ranks.size(); i++) {
		CommRank &r = c->ranks[i];
		snl_scene *s = r.scene;
		SNL_CUDA(cudaSetDevice(s->device));
		SNL_REQUIRE(s->nGpuNodes > 0, SNL_ERR_STATE, "rank %d: no geometry", r.rank);
		SNL_REQUIRE(!s->gVals[9], SNL_ERR_UNSUPPORTED, "gVals[9] supersampling is not available in snl_render_multi");
		// the register-budget variant of k_render is picked on this rank's first frame by the single-GPU path
		if(s->tune.choice < 0 || s->tuneW != W || s->tuneH != H) {
			SNL_CUDA(s->dImage.reserve((size_t)W * 3 * H));
			SNL_CUDA(cudaMemsetAsync(s->dCounters.ptr, 0, sizeof(Counters), s->stream));
			SNL_TRY(render_local(s, cam, W, H, nullptr, 0, s->dImage.ptr, (size_t)W * 3, r.rank));
			SNL_CUDA(cudaStreamSynchronize(s->stream));
			s->tuneW = W;
			s->tuneH = H;
		}
		if(c->ranks.size() == 1) {
			// one rank per process: open / wait for the frame slot first, so that the rank's device time (ev0 .. ev1: per-origin
			// records, its packets, its publish) does not count the time it was early for the slot
			CommCtl *ctl = reinterpret_cast<CommCtl *>(r.seg);
			if(r.rank == 0)
				SNL_CUDA(launch_frame_open(ctl, seq, c->world, s->stream));
			else
				SNL_CUDA(launch_wait_counter(&ctl->ready[slot], seq + 1u, &ctl->error, s->stream));
			s->launches++;
		}
		SNL_CUDA(cudaEventRecord(r.ev0, s->stream));
		SNL_CUDA(cudaMemsetAsync(s->dCounters.ptr, 0, sizeof(Counters), s->stream));
		SNL_TRY(prepare_frame(s, cam, W, H, r.rank, devScene[i], rayGen[i]));
	}
	// Phase 2: the frame itself, launches only. Rank 0 goes LAST and its device->host copies last of all: a copy into
	// pageable host memory blocks the calling thread until the stream gets there, and by then every other local rank
	// must already be enqueued (they wait on the device for rank 0 to open the slot, not on the host).
	for(size_t k = 0; k < c->ranks.size(); k++) {
		const size_t i = (k + 1) % c->ranks.size();
		CommRank &r = c->ranks[i];
		snl_scene *s = r.scene;
		const SceneDev &S = devScene[i];
		const RayGenParams &g = rayGen[i];
		SNL_CUDA(cudaSetDevice(s->device));
		CommCtl *ctl = reinterpret_cast<CommCtl *>(r.seg);
		unsigned char *frame = r.seg + kCommCtlBytes + (size_t)slot * c->frameBytes;
		isRoot = isRoot || r.rank == 0;
		if(c->ranks.size() > 1) {  // (a single local rank has done this in phase 1)
			if(r.rank == 0)
				SNL_CUDA(launch_frame_open(ctl, seq, c->world, s->stream));
			else  // the slot is open once rank 0 has read frame seq - 2 out of it and reset its counters
				SNL_CUDA(launch_wait_counter(&ctl->ready[slot], seq + 1u, &ctl->error, s->stream));
			s->launches++;
		}
		const int variant = std::max(0, s->tune.choice);
		if(r.sortPending) {  // this frame's packet order comes from the side stream
			SNL_CUDA(cudaStreamWaitEvent(s->stream, r.sortOut, 0));
			r.sortPending = false;
		}
		const int4 *list = r.dPackets[r.cur];
		const int nList = r.bandFirst[kCommBands];
		unsigned *cost = c->sortByCost ? r.dCost : nullptr;
		if(c->nBands == 1) {
			// one launch per rank; the frame is complete when every rank has published
			if(nList > 0) {
				SNL_CUDA(launch_render(S, s->originPre.set(), g, W, H, list, nList, frame, c->pitch, variant, s->stream, 0, nList, cost));
				s->launches++;
			}
		}
		else {
			SNL_CUDA(cudaEventRecord(s->forkEv, s->stream));
			for(int b = 0; b < c->nBands; b++) {
				cudaStream_t st = s->bandStream[b];
				SNL_CUDA(cudaStreamWaitEvent(st, s->forkEv, 0));
				const int first = r.bandFirst[b], count = r.bandFirst[b + 1] - first;
				if(count > 0) {
					SNL_CUDA(launch_render(S, s->originPre.set(), g, W, H, list, nList, frame, c->pitch, variant, st, first, count, cost));
					s->launches++;
				}
				SNL_CUDA(launch_band_signal(&ctl->bandDone[slot][b], st));
				s->launches++;
			}
			for(int b = 0; b < c->nBands; b++) {
				cudaStream_t st = s->bandStream[b];
				if(r.rank == 0 && bgr8_host) {
					const int r0 = b * c->rowsPerBand, r1 = std::min(py, (b + 1) * c->rowsPerBand);
					const int y0 = r0 * 16, y1 = std::min(H, r1 * 16);
					SNL_CUDA(launch_wait_counter(&ctl->bandDone[slot][b], (unsigned)c->world, &ctl->error, st));
					s->launches++;
					SNL_CUDA(cudaMemcpy2DAsync(bgr8_host + (size_t)y0 * hostPitch, hostPitch, frame + (size_t)y0 * c->pitch, c->pitch,
											   (size_t)W * 3, y1 - y0, cudaMemcpyDeviceToHost, st));
				}
				SNL_CUDA(cudaEventRecord(s->bandEv[b], st));
				SNL_CUDA(cudaStreamWaitEvent(s->stream, s->bandEv[b], 0));
			}
		}
		SNL_CUDA(launch_frame_publish(ctl, s->dCounters.ptr, seq, r.rank, s->stream));
		s->launches++;
		SNL_CUDA(cudaEventRecord(r.ev1, s->stream));
		if(r.rank == 0) {
			// every rank has published = every packet of the frame is stored and every counter is in
			SNL_CUDA(launch_wait_counter(&ctl->statsArrived[slot], (unsigned)c->world, &ctl->error, s->stream));
			s->launches++;
			if(c->nBands == 1 && bgr8_host)
				SNL_CUDA(cudaMemcpy2DAsync(bgr8_host, hostPitch, frame, c->pitch, (size_t)W * 3, H, cudaMemcpyDeviceToHost, s->stream));
			SNL_CUDA(cudaMemcpyAsync(c->hCtl, ctl, sizeof(CommCtl), cudaMemcpyDeviceToHost, s->stream));
		}
		if(cost && nList > 1) {
			// off the frame's critical path: the list for the next frame, every band longest packet first
			SNL_CUDA(cudaEventRecord(r.sortIn, s->stream));
			SNL_CUDA(cudaStreamWaitEvent(r.sortStream, r.sortIn, 0));
			SNL_CUDA(launch_sort_band_packets(list, cost, r.dPackets[r.cur ^ 1], r.dBandFirst, c->nBands, r.sortStream));
			SNL_CUDA(cudaEventRecord(r.sortOut, r.sortStream));
			s->launches++;
			r.cur ^= 1;
			r.sortPending = true;
		}
	}
	c->seq++;
	float ms = 0.0f;
	for(CommRank &r : c->ranks) {
		SNL_CUDA(cudaSetDevice(r.scene->device));
		SNL_CUDA(cudaStreamSynchronize(r.scene->stream));
		float m = 0.0f;
		if(cudaEventElapsedTime(&m, r.ev0, r.ev1) == cudaSuccess)
			ms = std::max(ms, m);
	}
	c->lastMs = ms;
	if(isRoot)
		SNL_REQUIRE(c->hCtl->error == 0, SNL_ERR_STATE,
					"snl_render_multi: a rank did not arrive within the time-out (every rank must call it once per frame)");
	if(stats) {
		memset(stats, 0, sizeof(*stats));
		Counters k;
		if(isRoot)
			memcpy(&k, c->hCtl->stats[slot], sizeof(k));  // summed over the ranks
		else {                                   // this rank's share
			snl_scene *s = c->ranks[0].scene;
			SNL_CUDA(cudaMemcpy(&k, s->dCounters.ptr, sizeof(k), cudaMemcpyDeviceToHost));
		}
		stats->primary = k.primary;
		stats->shadow = k.shadow;
		stats->reflection = k.reflection;
		stats->reflection_shadow = k.reflection_shadow;
		stats->transparency = k.transparency;
		stats->transparency_truncated = k.transp_truncated;
		stats->rays = k.primary + k.shadow + k.reflection + k.reflection_shadow + k.transparency;
		stats->node_visits = k.node_visits;
		stats->tri_tests = k.tri_tests;
		stats->gtri_tests = k.gtri_tests;
		stats->kernel_ms = ms;
	}
	return SNL_OK;
}

}  // extern "C"
