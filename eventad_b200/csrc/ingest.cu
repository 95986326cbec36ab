// eventad_b200/csrc/ingest.cu -- device-resident event ingest: a recording's event arrays live in HBM and
// windows are cut, filtered and down-sampled there, in the formats the rest of the path consumes.
//
// Reference code replaced (all CPU / numpy / numba / h5py there):
//   create_ms_to_idx                         /root/reference/scripts/downsample_events.py:36-42
//   downsample_events + _filter_events_resize /root/reference/scripts/downsample_events.py:91-122
//   _load_events                              /root/reference/src/dagr/data/dsec_utils.py:80-120
// Event arrays are int32 x, y, t (microseconds relative to t_offset, ascending) and int8 p on the device.
#include "common.cuh"

// ------------------------------------------------------------------------------------------------ ms -> first event index
// ms_to_idx[m] = number of events with t // 1000 < m (the cumulative count the reference builds from np.unique); t is
// ascending, so it is the lower bound of 1000 m.
__global__ void __launch_bounds__(256) ing_ms_to_idx_kernel(const int32_t* __restrict__ t, int64_t n, int64_t n_ms,
                                                            int64_t* __restrict__ out) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= n_ms) return;
    const int64_t key = m * 1000;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t)t[mid] < key) lo = mid + 1; else hi = mid;
    }
    out[m] = lo;
}

extern "C" int evad_ingest_ms_to_idx(const int32_t* t_us, int64_t n, int64_t n_ms, int64_t* ms_to_idx, void* stream) {
    if (!t_us || !ms_to_idx || n < 0 || n_ms < 0) return EVAD_ERR_ARG;
    if (n_ms == 0) return EVAD_OK;
    ing_ms_to_idx_kernel<<<evad_cdiv(n_ms, 256), 256, 0, (cudaStream_t)stream>>>(t_us, n, n_ms, ms_to_idx);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ window slice
// _load_events: events [idx0, idx0 + cnt) -> xy int16, t = time_window - tq + t_ev (int32), p = 2 p - 1 (int8), kept when
// t > 0 and y < height (height < 0: no crop); order preserved.
__global__ void __launch_bounds__(256) ing_slice_flag_kernel(const int32_t* __restrict__ y, const int32_t* __restrict__ t,
                                                             int64_t idx0, int64_t cnt, int64_t tq, int64_t time_window,
                                                             int height, uint32_t* __restrict__ flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cnt) return;
    const int32_t tr = (int32_t)(time_window - tq + (int64_t)t[idx0 + i]);
    flag[i] = (tr > 0 && (height < 0 || (int16_t)y[idx0 + i] < height)) ? 1u : 0u;
}

__global__ void __launch_bounds__(256) ing_slice_scatter_kernel(
    const int32_t* __restrict__ x, const int32_t* __restrict__ y, const int32_t* __restrict__ t, const int8_t* __restrict__ p,
    int64_t idx0, int64_t cnt, int64_t tq, int64_t time_window, const uint32_t* __restrict__ flag,
    const uint32_t* __restrict__ excl, int16_t* __restrict__ xy_out, int32_t* __restrict__ t_out, int8_t* __restrict__ p_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= cnt || !flag[i]) return;
    const uint32_t o = excl[i];
    xy_out[2 * o] = (int16_t)x[idx0 + i];
    xy_out[2 * o + 1] = (int16_t)y[idx0 + i];
    t_out[o] = (int32_t)(time_window - tq + (int64_t)t[idx0 + i]);
    p_out[o] = (int8_t)(2 * p[idx0 + i] - 1);
}

extern "C" size_t evad_ingest_slice_ws_bytes(int64_t cnt) {
    const size_t n = (size_t)(cnt > 0 ? cnt : 1);
    return evad_align(n * 4) * 2 + evad_align(evad_scan_tmp_elems(cnt) * 4) + 256;
}

extern "C" int evad_ingest_slice(const int32_t* x, const int32_t* y, const int32_t* t, const int8_t* p, int64_t idx0,
                                 int64_t cnt, int64_t tq, int64_t time_window, int height, int16_t* xy_out, int32_t* t_out,
                                 int8_t* p_out, int32_t* n_out, void* ws, size_t ws_bytes, void* stream) {
    if (!x || !y || !t || !p || !xy_out || !t_out || !p_out || !n_out || !ws || idx0 < 0 || cnt < 0) return EVAD_ERR_ARG;
    if (ws_bytes < evad_ingest_slice_ws_bytes(cnt)) return EVAD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (cnt == 0) {
        if (cudaMemsetAsync(n_out, 0, 4, st) != cudaSuccess) return EVAD_ERR_LAUNCH;
        return EVAD_OK;
    }
    EvadArena a(ws, ws_bytes);
    uint32_t* flag = a.take<uint32_t>(cnt);
    uint32_t* excl = a.take<uint32_t>(cnt);
    uint32_t* tmp = a.take<uint32_t>(evad_scan_tmp_elems(cnt));
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    ing_slice_flag_kernel<<<evad_cdiv(cnt, 256), 256, 0, st>>>(y, t, idx0, cnt, tq, time_window, height, flag);
    EVAD_CHECK_LAUNCH();
    EVAD_TRY(evad_exclusive_scan_u32(flag, excl, cnt, reinterpret_cast<uint32_t*>(n_out), tmp, st));
    ing_slice_scatter_kernel<<<evad_cdiv(cnt, 256), 256, 0, st>>>(x, y, t, p, idx0, cnt, tq, time_window, flag, excl, xy_out,
                                                                 t_out, p_out);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}

// ------------------------------------------------------------------------------------------------ down-sampling filter
// _filter_events_resize is a sequential loop over the events, but its state is per coarse cell: the events are stably
// sorted by cell (so a cell's events keep their order), one thread walks each cell's run with the cell's value in a
// register and marks the events that pass; survivors are compacted in their original order with x / fx, y / fy.
__global__ void __launch_bounds__(256) ing_ds_keys_kernel(const int32_t* __restrict__ x, const int32_t* __restrict__ y,
                                                          int64_t n, int fx, int fy, int out_w, int out_h,
                                                          uint32_t* __restrict__ keys, uint32_t* __restrict__ status) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int xl = x[i] / fx, yl = y[i] / fy;
    if (x[i] < 0 || y[i] < 0 || xl >= out_w || yl >= out_h) {
        atomicOr(status, EVAD_STATUS_EVENT_OOB);
        keys[i] = 0;
        return;
    }
    keys[i] = (uint32_t)(yl * out_w + xl);
}

__global__ void __launch_bounds__(128) ing_ds_walk_kernel(const uint32_t* __restrict__ skey, const uint32_t* __restrict__ sidx,
                                                          int64_t n, const int8_t* __restrict__ p, int fx, int fy,
                                                          float* __restrict__ change_map, uint32_t* __restrict__ flag) {
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const uint32_t key = skey[s];
    if (s > 0 && skey[s - 1] == key) return;                 // not the first event of its cell
    float v = change_map[key];
    const double inv = (double)(fx * fy);
    for (int64_t q = s; q < n && skey[q] == key; q++) {
        const uint32_t i = sidx[q];
        const double pv = (double)p[i];
        v = (float)((double)v + pv * 1.0 / inv);             // numba: float64 expression, float32 store
        uint32_t pass = 0u;
        if (fabsf(v) >= 1.f) {
            pass = 1u;
            v = (float)((double)v - pv);
        }
        flag[i] = pass;
    }
    change_map[key] = v;
}

__global__ void __launch_bounds__(256) ing_ds_scatter_kernel(const int32_t* __restrict__ x, const int32_t* __restrict__ y,
                                                             const int32_t* __restrict__ t, const int8_t* __restrict__ p,
                                                             int64_t n, int fx, int fy, const uint32_t* __restrict__ flag,
                                                             const uint32_t* __restrict__ excl, int32_t* __restrict__ x_out,
                                                             int32_t* __restrict__ y_out, int32_t* __restrict__ t_out,
                                                             int8_t* __restrict__ p_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flag[i]) return;
    const uint32_t o = excl[i];
    x_out[o] = x[i] / fx;                                    // (x / fx).astype(uint16): truncation of a non-negative quotient
    y_out[o] = y[i] / fy;
    t_out[o] = t[i];
    p_out[o] = p[i];
}

extern "C" size_t evad_ingest_downsample_ws_bytes(int64_t n) {
    const size_t m = (size_t)(n > 0 ? n : 1);
    return evad_align(m * 4) * 6 + evad_align(evad_sort_tmp_elems(n) * 4) + evad_align(evad_scan_tmp_elems(n) * 4) + 256;
}

// change_map [out_h, out_w] float32 is read and updated (the reference threads it through consecutive calls).
extern "C" int evad_ingest_downsample(const int32_t* x, const int32_t* y, const int32_t* t, const int8_t* p, int64_t n, int fx,
                                      int fy, int out_w, int out_h, float* change_map, int32_t* x_out, int32_t* y_out,
                                      int32_t* t_out, int8_t* p_out, int32_t* n_out, uint32_t* status, void* ws,
                                      size_t ws_bytes, void* stream) {
    if (!x || !y || !t || !p || !change_map || !x_out || !y_out || !t_out || !p_out || !n_out || !status || !ws || n < 0 ||
        fx < 1 || fy < 1 || out_w < 1 || out_h < 1)
        return EVAD_ERR_ARG;
    if ((int64_t)out_w * out_h >= (1ll << 31)) return EVAD_ERR_UNSUPPORTED;
    if (ws_bytes < evad_ingest_downsample_ws_bytes(n)) return EVAD_ERR_WORKSPACE;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        if (cudaMemsetAsync(n_out, 0, 4, st) != cudaSuccess) return EVAD_ERR_LAUNCH;
        return EVAD_OK;
    }
    EvadArena a(ws, ws_bytes);
    uint32_t* k0 = a.take<uint32_t>(n);
    uint32_t* v0 = a.take<uint32_t>(n);
    uint32_t* k1 = a.take<uint32_t>(n);
    uint32_t* v1 = a.take<uint32_t>(n);
    uint32_t* flag = a.take<uint32_t>(n);
    uint32_t* excl = a.take<uint32_t>(n);
    uint32_t* sort_tmp = a.take<uint32_t>(evad_sort_tmp_elems(n));
    uint32_t* scan_tmp = a.take<uint32_t>(evad_scan_tmp_elems(n));
    if (!a.ok) return EVAD_ERR_WORKSPACE;
    ing_ds_keys_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(x, y, n, fx, fy, out_w, out_h, k0, status);
    EVAD_CHECK_LAUNCH();
    int key_bits = 1;
    while ((1ll << key_bits) < (int64_t)out_w * out_h) key_bits++;
    uint32_t *skey, *sidx;
    EVAD_TRY(evad_sort_pairs_u32(k0, v0, k1, v1, 1, n, key_bits, sort_tmp, &skey, &sidx, st));
    ing_ds_walk_kernel<<<evad_cdiv(n, 128), 128, 0, st>>>(skey, sidx, n, p, fx, fy, change_map, flag);
    EVAD_CHECK_LAUNCH();
    EVAD_TRY(evad_exclusive_scan_u32(flag, excl, n, reinterpret_cast<uint32_t*>(n_out), scan_tmp, st));
    ing_ds_scatter_kernel<<<evad_cdiv(n, 256), 256, 0, st>>>(x, y, t, p, n, fx, fy, flag, excl, x_out, y_out, t_out, p_out);
    EVAD_CHECK_LAUNCH();
    return EVAD_OK;
}
