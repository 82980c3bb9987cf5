// One host batch over several GPUs from ONE process: frames are independent, so a batch of n
// stereo pairs is cut into contiguous shards, one per handle (= per GPU), and every shard runs the whole host-to-host
// path (sa_run_batch: pinned async H2D on a side stream, stage A, network slot, stage B/C, async D2H on another side
// stream) on its own host thread.  There is no collective and no peer traffic: each GPU reads its slice of the caller's
// input buffers and writes its slice of the output buffers.  The reference has nothing like it (one object = one GPU,
// cudaSetDevice in Initialize only, FastACVNet_plus/src/TRTFastACVNet_plus.cpp:45).
#include <thread>
#include <vector>

#include "common.h"

extern "C" {

int sa_shard_range(int n_total, int rank, int world, int* begin, int* end)
{
    if (world <= 0 || rank < 0 || rank >= world || n_total < 0 || !begin || !end) return SA_ERR_INVALID_ARG;
    const long long per = ((long long)n_total + world - 1) / world;      // ceil(n / world): the last shards are shorter, possibly empty
    long long b = (long long)rank * per;
    if (b > n_total) b = n_total;
    long long e = b + per;
    if (e > n_total) e = n_total;
    *begin = (int)b;
    *end = (int)e;
    return SA_OK;
}

int sa_run_batch_multi(sa_handle* const* handles, int n_handles, int n, const uint8_t* left, const uint8_t* right, int rectify, unsigned stub_seed,
                       float* pointcloud, float* disparity, uint8_t* rect_left, uint8_t* rect_right, uint8_t* heatmap, int* failed_handle)
{
    if (failed_handle) *failed_handle = -1;
    if (!handles || n_handles <= 0 || n < 0) return SA_ERR_INVALID_ARG;
    for (int g = 0; g < n_handles; g++)
        if (!handles[g]) return SA_ERR_INVALID_ARG;
    if (n == 0) return SA_OK;
    // all handles must describe the same camera and network: the shards are slices of the same arrays
    const sa_handle* h0 = handles[0];
    for (int g = 1; g < n_handles; g++) {
        const sa_handle* h = handles[g];
        if (h->W != h0->W || h->H != h0->H || h->w != h0->w || h->h != h0->h || h->cfg.model != h0->cfg.model || h->cfg.tensor_dtype != h0->cfg.tensor_dtype) {
            if (failed_handle) *failed_handle = g;
            return sa::fail(handles[g], SA_ERR_INVALID_ARG, "sa_run_batch_multi: handle %d differs from handle 0 in image size, network size, model or tensor type", g);
        }
    }
    const size_t img = (size_t)h0->W * h0->H * 3, npix = (size_t)h0->w * h0->h;
    std::vector<int> rc((size_t)n_handles, SA_OK);
    auto shard = [&](int g) {
        int b = 0, e = 0;
        sa_shard_range(n, g, n_handles, &b, &e);
        if (e <= b) return;
        sa_handle* h = handles[g];
        if (e - b > h->max_n) {
            rc[g] = sa::fail(h, SA_ERR_TOO_LARGE, "sa_run_batch_multi: shard %d holds %d pairs, the handle's max_batch is %d", g, e - b, h->max_n);
            return;
        }
        const long saved = h->opt_frame_base;
        h->opt_frame_base = saved + b;                   // the synthetic-disparity stub numbers frames globally: shards reproduce the one-GPU batch
        rc[g] = sa_run_batch(h, e - b, left + (size_t)b * img, right + (size_t)b * img, rectify, stub_seed, pointcloud ? pointcloud + (size_t)b * npix * 6 : nullptr,
                             disparity ? disparity + (size_t)b * npix : nullptr, rect_left ? rect_left + (size_t)b * img : nullptr,
                             rect_right ? rect_right + (size_t)b * img : nullptr, heatmap ? heatmap + (size_t)b * npix * 3 : nullptr);
        h->opt_frame_base = saved;
    };
    // one host thread per GPU (the calling thread takes shard 0); each sa_run_batch selects its own device
    std::vector<std::thread> workers;
    workers.reserve((size_t)n_handles);
    for (int g = 1; g < n_handles; g++) workers.emplace_back(shard, g);
    shard(0);
    for (auto& t : workers) t.join();
    // every shard has finished (sa_run_batch drains its streams on failure too): report the first failure
    for (int g = 0; g < n_handles; g++)
        if (rc[g] != SA_OK) {
            if (failed_handle) *failed_handle = g;
            return rc[g];
        }
    return SA_OK;
}

}  // extern "C"
