// reactor_types.h — reactor models of BatchReactorNet and the layout of their solution vectors.
#pragma once

namespace ctb {

//! Reactor models (CTB_REACTOR_* of include/ctb_clib.h) and the layout of their solution vectors:
//!   0 IdealGasConstPressureMoleReactor  y = [T, n_1..n_K]        (IdealGasConstPressureMoleReactor.cpp:22-30)
//!   1 IdealGasMoleReactor               y = [T, V, n_1..n_K]     (IdealGasMoleReactor.cpp:97-113)
//!   2 IdealGasReactor                   y = [m, V, T, Y_1..Y_K]  (IdealGasReactor.cpp:38-49)
//!   3 IdealGasConstPressureReactor      y = [m, T, Y_1..Y_K]     (IdealGasConstPressureReactor.cpp:36-46)
#ifndef CTB_RT_HD
#if defined(__CUDACC__)
#define CTB_RT_HD __host__ __device__
#else
#define CTB_RT_HD
#endif
#endif
CTB_RT_HD inline int reactorSIdx(int type) { return type == 0 ? 1 : type == 2 ? 3 : 2; }   // first species row
CTB_RT_HD inline int reactorTIdx(int type) { return type == 2 ? 2 : type == 3 ? 1 : 0; }   // temperature row
CTB_RT_HD inline int reactorNeq(int type, int K) { return K + reactorSIdx(type); }
CTB_RT_HD inline bool reactorIsMole(int type) { return type <= 1; }
CTB_RT_HD inline bool reactorIsConstP(int type) { return type == 0 || type == 3; }

} // namespace ctb
