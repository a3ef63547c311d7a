// TEST INFRASTRUCTURE - stand-in for <ceres/ceres.h>: the part of the Ceres API that the
// reference's bundle_adjuster.cpp / reprojection_factor.hpp touch, as a RECORDING problem with
// no solver (ceres::Solve leaves every parameter block as it is - the same deviation the cv2
// oracle and the product's BundleAdjuster document: Ceres is not in this image). The problem
// keeps the live residual blocks (cost function, pose block, point block) and parameter-block
// state so that the harness can read back what the reference's add_keyframe / remove_oldest_pose
// built. See oracle/build_ref.py.
#pragma once
#include <algorithm>
#include <cstddef>
#include <iostream>
#include <map>
#include <memory>
#include <queue>
#include <set>
#include <utility>
#include <vector>

namespace ceres {

class CostFunction {
  public:
    virtual ~CostFunction() {}
    virtual bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const = 0;
};
template <int kNumResiduals, int... Ns>
class SizedCostFunction : public CostFunction {};

class LossFunction {
  public:
    virtual ~LossFunction() {}
};

class LocalParameterization {
  public:
    virtual ~LocalParameterization() {}
};
class QuaternionParameterization : public LocalParameterization {};
class IdentityParameterization : public LocalParameterization {
  public:
    explicit IdentityParameterization(int) {}
};
class ProductParameterization : public LocalParameterization {
  public:
    ProductParameterization(LocalParameterization* a, LocalParameterization* b) : a_(a), b_(b) {}
  private:
    std::unique_ptr<LocalParameterization> a_, b_;      // Ceres takes ownership
};

struct ResidualBlock {
    CostFunction* cost;
    double* pose;
    double* point;
};
typedef ResidualBlock* ResidualBlockId;

class Problem;
// the problem made last (the harness picks up the one the adjuster's constructor builds)
inline Problem*& last_problem() {
    static Problem* p = nullptr;
    return p;
}

class Problem {
  public:
    struct Options {
        bool enable_fast_removal = false;
    };
    explicit Problem(const Options& o) : options_(o) { last_problem() = this; }
    ~Problem() {
        for (ResidualBlock* b : blocks_) { delete b->cost; delete b; }
    }
    ResidualBlockId AddResidualBlock(CostFunction* cost, LossFunction*, double* x0, double* x1) {
        ResidualBlock* b = new ResidualBlock{cost, x0, x1};
        blocks_.push_back(b);
        params_.insert(x0);
        params_.insert(x1);
        return b;
    }
    void RemoveResidualBlock(ResidualBlockId id) {
        auto it = std::find(blocks_.begin(), blocks_.end(), id);
        if (it == blocks_.end()) { ++errors; return; }
        delete (*it)->cost;
        delete *it;
        blocks_.erase(it);
    }
    void RemoveParameterBlock(double* p) {
        if (!params_.erase(p)) ++errors;
        constant_.erase(p);
        // Ceres also drops every residual block that still depends on the parameter block
        for (size_t i = 0; i < blocks_.size();) {
            if (blocks_[i]->pose == p || blocks_[i]->point == p) {
                delete blocks_[i]->cost;
                delete blocks_[i];
                blocks_.erase(blocks_.begin() + (std::ptrdiff_t)i);
                ++dropped_with_parameter;
            } else {
                ++i;
            }
        }
    }
    void SetParameterization(double* p, LocalParameterization*) { if (!params_.count(p)) ++errors; }
    void SetParameterBlockConstant(double* p) {
        if (!params_.count(p)) ++errors;
        constant_.insert(p);
    }
    // ---- read-back for the harness (not Ceres API)
    const std::vector<ResidualBlock*>& blocks() const { return blocks_; }
    bool is_constant(double* p) const { return constant_.count(p) != 0; }
    size_t num_parameter_blocks() const { return params_.size(); }
    int errors = 0;                    // API misuse Ceres would abort on
    int dropped_with_parameter = 0;

  private:
    Options options_;
    std::vector<ResidualBlock*> blocks_;
    std::set<double*> params_, constant_;
};

enum LinearSolverType { DENSE_QR, DENSE_SCHUR, SPARSE_SCHUR };
enum TerminationType { CONVERGENCE, NO_CONVERGENCE, FAILURE };

class Solver {
  public:
    struct Options {
        LinearSolverType linear_solver_type = DENSE_QR;
        bool minimizer_progress_to_stdout = false;
        double max_solver_time_in_seconds = 1e9;
        int num_threads = 1;
    };
    struct Summary {
        TerminationType termination_type = CONVERGENCE;
    };
};

// no solver in this image: parameter blocks stay as add_keyframe left them
inline void Solve(const Solver::Options&, Problem*, Solver::Summary* s) { s->termination_type = CONVERGENCE; }

}  // namespace ceres
