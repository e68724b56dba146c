// two_device_test.cc — two orbfe handles on two devices, driven from two std::threads of ONE process, at the shape whose
// quadtree needs more than 48 KB of dynamic shared memory (1920x1080, 4000 features).  This is how the reference uses its
// extractors (one thread per eye, Frame.cc:119-122) mapped onto one GPU per thread.  Both devices must give identical
// results for identical frames.  Exit 77: fewer than two devices.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "orbfe.h"

extern "C" int orbfe_synth_frame(uint64_t seed, int w, int h, int view, int shift, uint8_t* out);

struct Result {
    int status = 0;
    std::vector<int> n, mono;
    std::vector<orbfe_keypoint> kps;
    std::vector<uint8_t> desc;
};

static void run(int device, const std::vector<uint8_t>& frames, int nframes, int W, int H, Result* out) {
    orbfe_params p;
    p.nfeatures = 4000; p.scale_factor = 1.2f; p.nlevels = 8; p.ini_th_fast = 20; p.min_th_fast = 7;
    p.device = device; p.max_width = W; p.max_height = H; p.max_batch = nframes;
    orbfe_handle* h = nullptr;
    out->status = orbfe_create(&p, &h);
    if (out->status != ORBFE_OK) return;
    const int cap = orbfe_max_keypoints(h);
    out->n.assign(nframes, 0);
    out->mono.assign(nframes, 0);
    out->kps.assign((size_t)nframes * cap, orbfe_keypoint());
    out->desc.assign((size_t)nframes * cap * 32, 0);
    for (int rep = 0; rep < 3 && out->status >= 0; ++rep)       // eager, graph capture, graph replay
        out->status = orbfe_extract_batch(h, frames.data(), nframes, W, H, W, (ptrdiff_t)W * H, 0, 0, out->kps.data(), out->desc.data(), cap,
                                          out->n.data(), out->mono.data());
    if (out->status >= 0) {                                        // and the one-image drop-in call
        std::vector<orbfe_keypoint> k(cap);
        std::vector<uint8_t> d((size_t)cap * 32);
        int n1 = 0;
        const int m1 = orbfe_extract(h, frames.data(), W, H, W, 0, 0, k.data(), d.data(), cap, &n1);
        if (m1 < 0) out->status = m1;
        else if (n1 != out->n[0] || memcmp(k.data(), out->kps.data(), sizeof(orbfe_keypoint) * (size_t)n1)) out->status = -100;
    }
    if (out->status < 0) fprintf(stderr, "device %d: %s %s\n", device, orbfe_strerror(out->status), orbfe_last_cuda_error());
    // trim to the valid rows so the comparison below does not read unspecified tails
    for (int b = 0; b < nframes; ++b) {
        memset(out->kps.data() + (size_t)b * cap + out->n[b], 0, sizeof(orbfe_keypoint) * (size_t)(cap - out->n[b]));
        memset(out->desc.data() + ((size_t)b * cap + out->n[b]) * 32, 0, (size_t)(cap - out->n[b]) * 32);
    }
    orbfe_destroy(h);
}

int main() {
    const int W = 1920, H = 1080, N = 3;
    std::vector<uint8_t> frames((size_t)N * W * H);
    for (int i = 0; i < N; ++i)
        if (orbfe_synth_frame(1400 + i, W, H, 0, 0, frames.data() + (size_t)i * W * H) != 0) return 2;
    Result probe;
    {
        orbfe_params p = {1000, 1.2f, 8, 20, 7, 1, 752, 480, 1};
        orbfe_handle* h = nullptr;
        const int st = orbfe_create(&p, &h);
        if (st == ORBFE_ERR_BAD_ARGUMENT || st == ORBFE_ERR_NO_DEVICE) { printf("fewer than two devices\n"); return 77; }
        orbfe_destroy(h);
    }
    Result r[2];
    std::thread t0(run, 0, std::cref(frames), N, W, H, &r[0]), t1(run, 1, std::cref(frames), N, W, H, &r[1]);
    t0.join();
    t1.join();
    if (r[0].status < 0 || r[1].status < 0) return 1;
    if (r[0].n != r[1].n || r[0].mono != r[1].mono || memcmp(r[0].kps.data(), r[1].kps.data(), r[0].kps.size() * sizeof(orbfe_keypoint)) ||
        r[0].desc != r[1].desc) {
        printf("results differ between the devices\n");
        return 1;
    }
    printf("identical on 2 devices: %d + %d + %d keypoints\n", r[0].n[0], r[0].n[1], r[0].n[2]);
    return r[0].n[0] >= 4000 ? 0 : 1;
}
