// TEMPORARY stub (replaced by the BDF kernels)
#include "reactor.h"
namespace ctb {
long long reactorHistDoubles(int neq) { return 8LL * neq + 32; }
void launchAdvance(const DeviceMech&, const ReactorBatch&, cudaStream_t) { throw CtbError("advance", "not implemented"); }
void launchReactorInit(const DeviceMech&, long long, const double*, const double*, const double*, long long, double*, double*, double*, cudaStream_t) { throw CtbError("reactorInit", "not implemented"); }
void launchReactorContents(const DeviceMech&, long long, const double*, double*, double*, long long, cudaStream_t) { throw CtbError("reactorContents", "not implemented"); }
}
