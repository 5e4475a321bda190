// oracle/shim/ceres/solver.h — empty stand-in (only BundleAdjuster.h's solver members use it, and those are not compiled here).
#ifndef BA_SHIM_CERES_SOLVER_H
#define BA_SHIM_CERES_SOLVER_H
#endif
