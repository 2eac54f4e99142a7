// fes_hess.cu -- the second-order tangent kernel (fes_hess.cuh) for the six (model kind, fatigue) pairs, with and
// without the force factor as a direction.
#include "fes_hess.cuh"

namespace fes {

template <int KIND, bool FAT, bool FC>
static int launch_hess_one(const HessArgs &a, cudaStream_t s) {
    const int64_t total = (int64_t)a.ns * a.npairs * a.B;
    // 148 SMs x 16 resident blocks of 128 threads cover the device; larger problems stride over the grid
    const int64_t blocks = std::min<int64_t>((total + 127) / 128, (int64_t)148 * 16 * 8);
    fes_hess_kernel<KIND, FAT, FC><<<(unsigned)blocks, 128, (size_t)(2 * a.S + 1) * sizeof(double), s>>>(a);
    return (int)cudaGetLastError();
}

template <int KIND, bool FAT>
static int launch_hess_kf(const HessArgs &a, cudaStream_t s) {
    return a.factor ? launch_hess_one<KIND, FAT, true>(a, s) : launch_hess_one<KIND, FAT, false>(a, s);
}

int launch_hess(int kind, int fat, const HessArgs &a, cudaStream_t s) {
    switch (kind * 2 + (fat ? 1 : 0)) {
        case 0: return launch_hess_kf<DING2003, false>(a, s);
        case 1: return launch_hess_kf<DING2003, true>(a, s);
        case 2: return launch_hess_kf<DING2007, false>(a, s);
        case 3: return launch_hess_kf<DING2007, true>(a, s);
        case 4: return launch_hess_kf<HMED2018, false>(a, s);
        case 5: return launch_hess_kf<HMED2018, true>(a, s);
    }
    return -2;
}

}  // namespace fes
