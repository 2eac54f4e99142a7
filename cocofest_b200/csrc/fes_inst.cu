// fes_inst.cu -- explicit instantiation of the launchers for ONE (model kind, fatigue) pair.
// Compiled six times (-DFES_KIND=0..2 -DFES_FAT=0..1) so that the translation units build in parallel.
#include "fes_kernels.cuh"
#include "fes_ident.cuh"

#ifndef FES_KIND
#error "compile with -DFES_KIND=<0|1|2> -DFES_FAT=<0|1>"
#endif

namespace fes {
template int launch_ivp_kf<double, FES_KIND, (FES_FAT != 0)>(int, const IvpArgs<double> &, cudaStream_t);
template int launch_shoot_kf<double, FES_KIND, (FES_FAT != 0)>(int, const ShootArgs<double> &, cudaStream_t);
template int launch_ivp_kf<float, FES_KIND, (FES_FAT != 0)>(int, const IvpArgs<float> &, cudaStream_t);
template int launch_shoot_kf<float, FES_KIND, (FES_FAT != 0)>(int, const ShootArgs<float> &, cudaStream_t);
#if FES_FAT == 0
template int launch_ident_k<double, FES_KIND>(const IdentArgs<double> &, cudaStream_t);
template int launch_ident_k<float, FES_KIND>(const IdentArgs<float> &, cudaStream_t);
#endif
}  // namespace fes
