// synth.cu -- counter-based synthetic clouds generated directly in HBM for bench.py and the
// full-size property tests.  Point i is a pure function of (kind, n_total, seed, i): u(i,c) is
// splitmix64(seed*GOLDEN + stream*K + i*16 + c) / 2^53.  Densities are held at the levels of
// BASELINE.json's configs by scaling the extent with n_total (strip: 5000 pts per metre of corridor;
// forest: 500 pts/m^2 of plot; city: 555 pts/m^2 of block = 200 M points on 600 m x 600 m).
#include "common.cuh"

namespace {

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

struct Rng {
    unsigned long long base;
    __device__ Rng(unsigned long long seed, unsigned long long stream, unsigned long long i)
        : base(seed * 0x9E3779B97F4A7C15ull + stream * 0xD1B54A32D192ED03ull + i * 16ull)
    {
    }
    __device__ double u(int c) const
    {
        return (double)(splitmix64(base + (unsigned long long)c) >> 11) * (1.0 / 9007199254740992.0);
    }
    __device__ double gauss(int c) const // Box-Muller on u(c), u(c+1)
    {
        double a = u(c), b = u(c + 1);
        if (a < 1e-300) a = 1e-300;
        return sqrt(-2.0 * log(a)) * cos(6.283185307179586 * b);
    }
};

__device__ void tree_point(const Rng& r, unsigned long long seed, unsigned long long ntree, double side, bool crown,
                           double ground_amp, double& x, double& y, double& z)
{
    unsigned long long t = (unsigned long long)(r.u(1) * (double)ntree);
    if (t >= ntree) t = ntree - 1;
    Rng tr(seed, 7, t);
    const double tx = tr.u(0) * side, ty = tr.u(1) * side;
    const double rad = 0.1 + 0.3 * tr.u(2), hgt = 8.0 + 17.0 * tr.u(3), sig = 1.0 + tr.u(4);
    const double gz = ground_amp * sin(tx / 15.0) * cos(ty / 20.0);
    if (!crown) { // trunk: uniform on the cylinder surface
        const double a = 6.283185307179586 * r.u(2);
        x = tx + rad * cos(a);
        y = ty + rad * sin(a);
        z = gz + hgt * r.u(3);
    } else { // crown: isotropic Gaussian blob at the trunk top
        x = tx + sig * r.gauss(2);
        y = ty + sig * r.gauss(4);
        z = gz + hgt + sig * r.gauss(6);
    }
}

__global__ void __launch_bounds__(256) synth_kernel(int kind, unsigned long long n_total, unsigned long long first,
                                                    unsigned long long n, unsigned long long seed, int centred,
                                                    char* __restrict__ out)
{
    unsigned long long li = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
    if (li >= n) return;
    const unsigned long long i = first + li;
    Rng r(seed, 0, i);
    double x = 0, y = 0, z = 0;
    const double nt = (double)n_total;
    if (kind == 0) { // uniform_cube, side 10 m
        x = 10.0 * r.u(0), y = 10.0 * r.u(1), z = 10.0 * r.u(2);
    } else if (kind == 1) { // strip: 30 m wide mobile-mapping corridor, 2000 m per 10 M points
        const double L = fmax(nt * 2e-4, 20.0);
        const double m = r.u(0);
        if (m < 0.55) { // ground
            x = L * r.u(1), y = -15.0 + 30.0 * r.u(2), z = 0.02 * r.gauss(3);
        } else if (m < 0.85) { // two facades
            x = L * r.u(1);
            y = (r.u(2) < 0.5 ? -15.0 : 15.0) + 0.02 * r.gauss(3);
            z = 12.0 * r.u(5);
        } else if (m < 0.95) { // poles, one per 5 m
            unsigned long long npole = (unsigned long long)fmax(L / 5.0, 1.0);
            unsigned long long p = (unsigned long long)(r.u(1) * (double)npole);
            Rng pr(seed, 5, p);
            const double px = pr.u(0) * L, py = -14.0 + 28.0 * pr.u(1);
            const double a = 6.283185307179586 * r.u(2);
            x = px + 0.15 * cos(a), y = py + 0.15 * sin(a), z = 8.0 * r.u(3);
        } else { // clutter
            x = L * r.u(1), y = -15.0 + 30.0 * r.u(2), z = 12.0 * r.u(3);
        }
    } else if (kind == 2) { // forest plot, 200 m x 200 m per 20 M points
        const double side = fmax(200.0 * sqrt(nt / 2e7), 10.0);
        const unsigned long long ntree = (unsigned long long)fmax(1500.0 * nt / 2e7, 1.0);
        const double m = r.u(0);
        if (m < 0.35) {
            x = side * r.u(1), y = side * r.u(2);
            z = 0.5 * sin(x / 15.0) * cos(y / 20.0) + 0.03 * r.gauss(3);
        } else {
            tree_point(r, seed, ntree, side, m >= 0.5, 0.5, x, y, z);
        }
    } else if (kind == 3) { // city block, 600 m x 600 m per 200 M points
        const double side = fmax(600.0 * sqrt(nt / 2e8), 20.0);
        const double m = r.u(0);
        if (m < 0.40) { // ground
            x = side * r.u(1), y = side * r.u(2), z = 0.02 * r.gauss(3);
        } else if (m < 0.85) { // axis-aligned buildings: 4 walls + roof, picked by area
            const unsigned long long nbox = (unsigned long long)fmax(120.0 * nt / 2e8, 1.0);
            unsigned long long b = (unsigned long long)(r.u(1) * (double)nbox);
            if (b >= nbox) b = nbox - 1;
            Rng br(seed, 9, b);
            const double fx = fmin(15.0 + 45.0 * br.u(0), side * 0.5), fy = fmin(15.0 + 45.0 * br.u(1), side * 0.5);
            const double hh = 6.0 + 54.0 * br.u(2);
            const double bx = br.u(3) * (side - fx), by = br.u(4) * (side - fy);
            const double a1 = fx * hh, a2 = fy * hh, a3 = fx * fy, at = 2 * a1 + 2 * a2 + a3;
            const double f = r.u(2) * at;
            const double s = r.u(3), t = r.u(4), off = 0.01 * r.gauss(5);
            if (f < a1) {
                x = bx + s * fx, y = by + off, z = t * hh;
            } else if (f < 2 * a1) {
                x = bx + s * fx, y = by + fy + off, z = t * hh;
            } else if (f < 2 * a1 + a2) {
                x = bx + off, y = by + s * fy, z = t * hh;
            } else if (f < 2 * a1 + 2 * a2) {
                x = bx + fx + off, y = by + s * fy, z = t * hh;
            } else {
                x = bx + s * fx, y = by + t * fy, z = hh + off;
            }
        } else if (m < 0.95) { // trees
            const unsigned long long ntree = (unsigned long long)fmax(3000.0 * nt / 2e8, 1.0);
            tree_point(r, seed, ntree, side, m >= 0.88, 0.0, x, y, z);
        } else { // clutter
            x = side * r.u(1), y = side * r.u(2), z = 60.0 * r.u(3);
        }
    } else { // lattice with deliberate distance ties (spacing 0.25 m, every 16th point a duplicate)
        unsigned long long mm = (unsigned long long)ceil(cbrt(nt));
        if (mm < 1) mm = 1;
        unsigned long long j = (i % 16 == 15) ? i - 1 : i;
        x = 0.25 * (double)(j % mm), y = 0.25 * (double)((j / mm) % mm), z = 0.25 * (double)(j / (mm * mm));
    }
    if (!centred) { // UTM-scale world offset (magnitudes of file_helper.cpp:586)
        x += 483189.100569;
        y += 3371741.645173;
        z += 430.094726;
    }
    char* o = out + li * 48;
    double* od = reinterpret_cast<double*>(o);
    od[0] = x;
    od[1] = y;
    od[2] = z;
    od[3] = 1.0;
    *reinterpret_cast<uint32_t*>(o + 32) = (uint32_t)(splitmix64(i ^ (seed << 32)) & 0xffffffull);
    *reinterpret_cast<uint32_t*>(o + 36) = (uint32_t)(i % 10ull);
    *reinterpret_cast<unsigned long long*>(o + 40) = 0ull;
}

} // namespace

extern "C" int pcp_synth_cloud(pcp_ctx* ctx, int kind, uint64_t n_total, uint64_t first, uint64_t n, uint64_t seed,
                               int centred, void* out_pts)
{
    if (!ctx || kind < 0 || kind > 4 || (n && !out_pts) || first + n > n_total)
        return pcp_fail(ctx, PCP_ERR_INVALID, "pcp_synth_cloud: bad argument");
    if (n == 0) return PCP_OK;
    std::lock_guard<std::recursive_mutex> lock(ctx->mu);
    PCP_CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    OutStage o;
    PCP_TRY(o.open(ctx, out_pts, n * 48));
    synth_kernel<<<pcp_blocks_for(n, 256), 256, 0, ctx->stream>>>(kind, n_total, first, n, seed, centred,
                                                                 static_cast<char*>(o.dptr));
    PCP_LAUNCH_CHECK(ctx);
    PCP_TRY(o.commit());
    PCP_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return PCP_OK;
}
