// oracle/shim/ceres/ceres.h — just enough of <ceres/ceres.h> for the reference's cost-functor headers to compile (the factory
// SnavelyReprojectionError::Create, SnavelyReprojectionError.h:52-55, names CostFunction and AutoDiffCostFunction).
// TEST INFRASTRUCTURE ONLY; no solver here — the trust-region loop is restated in oracle/ba_oracle.cpp.
#ifndef BA_SHIM_CERES_CERES_H
#define BA_SHIM_CERES_CERES_H
#include "rotation.h"

namespace ceres {
class CostFunction {
public:
    virtual ~CostFunction() = default;
};
template <typename Functor, int kNumResiduals, int... Ns>
class AutoDiffCostFunction : public CostFunction {
public:
    explicit AutoDiffCostFunction(Functor* f) : functor_(f) {}
    ~AutoDiffCostFunction() override { delete functor_; }
    const Functor& functor() const { return *functor_; }
private:
    Functor* functor_;
};
}  // namespace ceres
#endif
