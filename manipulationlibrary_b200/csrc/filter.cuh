// filter.cuh -- the region filter, cython_filter_new/filter_cython_new.pyx:9-93, one CTA per entity.
//
//   seed     pyx:18-31  first point whose FLOAT32 distance to the per-axis median is strictly smallest
//   radius   pyx:35,55-59  float32(median_distance), compared against float64 distances with <=
//   CLOSURE mode           the fixed point of pyx:61-83: every point reachable from the seed through hops of
//                          length <= radius (frontier expansion; same set as the oracle's closure_filter)
//   REFERENCE_SEEDED mode  pyx:61-83 replayed pass by pass: scan the remaining points against the pivot in
//                          order, append the hits in scan order, draw the next pivot with CPython's
//                          random.choice (MT19937, _randbelow by rejection), stop after 30 dry passes
// Output per entity: flist = positions of the accepted points in output order WITHOUT the leading seed
// copy (the reference returns the seed twice: dicti[0] and again when pass 1 accepts it), n_list, seed_pos.
#pragma once
#include "common.cuh"

#define E2_FILTER_THREADS 256

// largest float64 x with sqrt(x) <= r  (so that  sqrt(d2) <= r  <=>  d2 <= x  exactly)
__device__ inline double sq_threshold(double r)
{
    if (!(r >= 0.0)) return -1.0;
    double t = r * r;
    while (sqrt(t) > r) t = __longlong_as_double(__double_as_longlong(t) - 1);
    for (;;) {
        double u = __longlong_as_double(__double_as_longlong(t) + 1);
        if (t == 0.0) u = 4.9406564584124654e-324;
        if (sqrt(u) <= r && u > t) t = u; else break;
    }
    return t;
}

struct Mt19937 {     // CPython's _random: init_by_array([seed]) as random.seed(int) does for 0 <= seed < 2^32
    unsigned mt[624];
    int idx;
    __device__ void init_genrand(unsigned s)
    {
        mt[0] = s;
        for (int i = 1; i < 624; ++i) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (unsigned)i;
        idx = 624;
    }
    __device__ void seed(unsigned key)
    {
        init_genrand(19650218u);
        int i = 1, j = 0;
        for (int k = 624; k; --k) {
            mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1664525u)) + key + (unsigned)j;
            ++i; ++j;
            if (i >= 624) { mt[0] = mt[623]; i = 1; }
            if (j >= 1) j = 0;
        }
        for (int k = 623; k; --k) {
            mt[i] = (mt[i] ^ ((mt[i - 1] ^ (mt[i - 1] >> 30)) * 1566083941u)) - (unsigned)i;
            ++i;
            if (i >= 624) { mt[0] = mt[623]; i = 1; }
        }
        mt[0] = 0x80000000u;
        idx = 624;
    }
    __device__ unsigned next()
    {
        if (idx >= 624) {
            int kk = 0;
            for (; kk < 624 - 397; ++kk) {
                unsigned y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
                mt[kk] = mt[kk + 397] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            for (; kk < 623; ++kk) {
                unsigned y = (mt[kk] & 0x80000000u) | (mt[kk + 1] & 0x7fffffffu);
                mt[kk] = mt[kk + (397 - 624)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            unsigned y = (mt[623] & 0x80000000u) | (mt[0] & 0x7fffffffu);
            mt[623] = mt[396] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            idx = 0;
        }
        unsigned y = mt[idx++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    // random.Random._randbelow_with_getrandbits(n), n >= 1
    __device__ unsigned randbelow(unsigned n)
    {
        int k = 32 - __clz(n);            // n.bit_length()
        unsigned r = next() >> (32 - k);
        while (r >= n) r = next() >> (32 - k);
        return r;
    }
};

struct FilterSmem {
    float best_d[E2_FILTER_THREADS / 32];
    int best_i[E2_FILTER_THREADS / 32];
    int seed;
    unsigned n_new, n_front;
    unsigned warp_cnt[E2_FILTER_THREADS / 32];
    unsigned n_out, n_keep, any_hit;
    double r2max;
    double pv[3];
    double fx[E2_FILTER_THREADS], fy[E2_FILTER_THREADS], fz[E2_FILTER_THREADS];   // frontier tile (CLOSURE)
};

// grid F*K CTAs. flist / front_a / front_b / member are indexed like the stored points (f*cap + off + i).
__global__ void __launch_bounds__(E2_FILTER_THREADS) region_filter_kernel(Dev d)
{
    pdl_enter();
    __shared__ FilterSmem sm;
    __shared__ Mt19937 rng;
    const int fe = blockIdx.x, f = fe / d.K, e = fe - f * d.K;
    const int off = d.ent_off[fe], n = d.ent_raw[fe];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (d.radius_mode == E2SEC_RADIUS_BOUNDED && d.amb[fe] != 1) return;  // decided by cell_filter_kernel (first or second visit)
    if (off < 0) {
        if (tid == 0) { d.n_list[fe] = 0; d.seed_pos[fe] = -1; }
        return;
    }
    const size_t pb = (size_t)f * d.cap + off;
    unsigned *flist = d.flist + pb;
    if (!entity_filtered(d, fe, n)) {                         // :1543-1546 the raw cloud is used
        for (int i = tid; i < n; i += E2_FILTER_THREADS) flist[i] = (unsigned)i;
        if (tid == 0) { d.n_list[fe] = n; d.seed_pos[fe] = -1; }
        return;
    }
    const double *X = d.px + pb, *Y = d.py + pb, *Z = d.pz + pb;
    const double cx = d.med[(size_t)fe * 3], cy = d.med[(size_t)fe * 3 + 1], cz = d.med[(size_t)fe * 3 + 2];
    // ---- seed: pyx:18-31 ------------------------------------------------------------------------
    float bd = 50000.0f;
    int bi = 0x7fffffff;
    for (int i = tid; i < n; i += E2_FILTER_THREADS) {
        float dt = (float)sqrt(dist2(X[i], Y[i], Z[i], cx, cy, cz));
        if (dt < bd) { bd = dt; bi = i; }     // strided scan keeps each thread's first minimum
    }
    for (int o = 16; o; o >>= 1) {
        float od = __shfl_xor_sync(0xffffffffu, bd, o);
        int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; }
    }
    if (lane == 0) { sm.best_d[warp] = bd; sm.best_i[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < E2_FILTER_THREADS / 32; ++w)
            if (sm.best_d[w] < bd || (sm.best_d[w] == bd && sm.best_i[w] < bi)) { bd = sm.best_d[w]; bi = sm.best_i[w]; }
        sm.seed = (bi == 0x7fffffff) ? 0 : bi;
        const double r = (double)(float)d.radius[fe];      // cdef float mean_distance = median_distance
        sm.r2max = sq_threshold(r);
        d.seed_pos[fe] = sm.seed;
    }
    __syncthreads();
    const int seed = sm.seed;
    const double r2max = sm.r2max;

    if (d.filter_mode == E2SEC_FILTER_CLOSURE) {
        unsigned char *member = d.member + pb;
        unsigned *fa = d.front_a + pb, *fb = d.front_b + pb;
        for (int i = tid; i < n; i += E2_FILTER_THREADS) member[i] = 0;
        if (tid == 0) { fa[0] = (unsigned)seed; sm.n_front = 1; sm.n_new = 0; }
        __syncthreads();
        for (;;) {
            const unsigned nf = sm.n_front;
            // the frontier is staged tile by tile in shared memory; every thread sweeps its not-yet-member points over it
            for (unsigned tile = 0; tile < nf; tile += E2_FILTER_THREADS) {
                __syncthreads();
                if (tile + tid < nf) {
                    const unsigned j = fa[tile + tid];
                    sm.fx[tid] = X[j]; sm.fy[tid] = Y[j]; sm.fz[tid] = Z[j];
                }
                __syncthreads();
                const unsigned cnt = min((unsigned)E2_FILTER_THREADS, nf - tile);
                for (int i = tid; i < n; i += E2_FILTER_THREADS) {
                    if (member[i]) continue;
                    const double x = X[i], y = Y[i], z = Z[i];
                    for (unsigned q = 0; q < cnt; ++q) {
                        if (dist2(x, y, z, sm.fx[q], sm.fy[q], sm.fz[q]) <= r2max) {
                            member[i] = 1;
                            fb[atomicAdd(&sm.n_new, 1u)] = (unsigned)i;
                            break;
                        }
                    }
                }
            }
            __syncthreads();
            const unsigned nn = sm.n_new;
            __syncthreads();
            if (nn == 0) break;
            if (tid == 0) { sm.n_front = nn; sm.n_new = 0; }
            unsigned *t = fa; fa = fb; fb = t;
            __syncthreads();
        }
        // stable compaction of the members (raster order)
        unsigned run = 0;
        for (int base = 0; base < n; base += E2_FILTER_THREADS) {
            const int i = base + tid;
            const bool m = (i < n) && member[i];
            const unsigned bal = __ballot_sync(0xffffffffu, m);
            if (lane == 0) sm.warp_cnt[warp] = __popc(bal);
            __syncthreads();
            unsigned before = run;
            for (int w = 0; w < warp; ++w) before += sm.warp_cnt[w];
            unsigned tot = 0;
            for (int w = 0; w < E2_FILTER_THREADS / 32; ++w) tot += sm.warp_cnt[w];
            if (m) flist[before + __popc(bal & ((1u << lane) - 1u))] = (unsigned)i;
            run += tot;
            __syncthreads();
        }
        if (tid == 0) d.n_list[fe] = (int)run;
        return;
    }

    // ---- REFERENCE_SEEDED: pyx:61-83 --------------------------------------------------------------
    unsigned *rem = d.front_a + pb, *nxt = d.front_b + pb;
    for (int i = tid; i < n; i += E2_FILTER_THREADS) rem[i] = (unsigned)i;
    if (tid == 0) {
        const unsigned s = d.filter_seed + 1000003u * (unsigned)d.frame_ids[f] + 7919u * (unsigned)(e - d.ent_shift);      // label_index as the caller counts it
        rng.seed(s);
        sm.n_out = 0;
        sm.pv[0] = X[seed]; sm.pv[1] = Y[seed]; sm.pv[2] = Z[seed];
    }
    __syncthreads();
    int n_rem = n, abort_cnt = 0;
    while (abort_cnt < 30) {
        const double qx = sm.pv[0], qy = sm.pv[1], qz = sm.pv[2];
        unsigned n_out = sm.n_out, n_keep = 0;
        const unsigned out0 = n_out;
        for (int base = 0; base < n_rem; base += E2_FILTER_THREADS) {
            const int k = base + tid;
            unsigned i = 0;
            bool hit = false, live = k < n_rem;
            if (live) {
                i = rem[k];
                hit = dist2(X[i], Y[i], Z[i], qx, qy, qz) <= r2max;
            }
            const unsigned bh = __ballot_sync(0xffffffffu, hit), bl = __ballot_sync(0xffffffffu, live);
            if (lane == 0) sm.warp_cnt[warp] = (unsigned)__popc(bh) | ((unsigned)__popc(bl) << 16);
            __syncthreads();
            unsigned hb = 0, lb = 0, ht = 0, ltot = 0;
            for (int w = 0; w < E2_FILTER_THREADS / 32; ++w) {
                const unsigned v = sm.warp_cnt[w];
                if (w < warp) { hb += v & 0xffffu; lb += v >> 16; }
                ht += v & 0xffffu; ltot += v >> 16;
            }
            const unsigned lm = (1u << lane) - 1u;
            if (hit) flist[n_out + hb + __popc(bh & lm)] = i;
            else if (live) nxt[n_keep + (lb - hb) + __popc((bl & ~bh) & lm)] = i;
            n_out += ht;
            n_keep += ltot - ht;
            __syncthreads();
        }
        abort_cnt = (n_out == out0) ? abort_cnt + 1 : 0;       // pyx:80-83
        n_rem = (int)n_keep;
        unsigned *t = rem; rem = nxt; nxt = t;
        if (n_rem == 0) __syncthreads();                       // a dry pass ran no barrier: every thread has read sm.pv / sm.n_out
        if (tid == 0) {
            sm.n_out = n_out;
            const unsigned c = rng.randbelow(n_out + 1u);      // random.choice(list(dicti.values())), pyx:78
            const unsigned j = (c == 0) ? (unsigned)seed : flist[c - 1];
            sm.pv[0] = X[j]; sm.pv[1] = Y[j]; sm.pv[2] = Z[j];
        }
        __syncthreads();
    }
    if (tid == 0) d.n_list[fe] = (int)sm.n_out;
}
