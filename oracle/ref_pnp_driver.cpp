// Checker driver (TEST INFRASTRUCTURE ONLY -- see oracle/README.md): evaluates the REFERENCE's own uncertainty-PnP cost
// functor, ReprojectionErrorArray::operator() (lib/utils/extend_utils/src/uncertainty_pnp.cpp:7-33), and its Jacobian by
// the reference's own automatic differentiation type (ceres::Jet from the vendored, header-only ceres/jet.h +
// ceres/rotation.h + Eigen under lib/utils/extend_utils/include).
//
// The functor is NOT copied into this repository: oracle/build.py cuts the struct out of the reference source where it
// lies into oracle/_ref/pnp_functor.inc at build time (git-ignored), without the AutoDiffCostFunction factory --
// <ceres/ceres.h> needs glog headers, which are absent, and libceres.so.1.14.0 cannot be linked (its dependencies glog /
// SuiteSparse / OpenBLAS are absent).  So the reference's COST is pinned by this file; its SOLVER (Ceres LM, 50
// iterations, function tolerance 1e-6) is not.
#include <ceres/jet.h>
#include <ceres/rotation.h>

using namespace std;            // as in the reference translation unit
#include "pnp_functor.inc"      // struct ReprojectionErrorArray { ctor; template <T> operator(); members }

extern "C" {

// residuals [pn,2] and Jacobian [pn,2,6] (d residual / d pose, pose = angle-axis[3] + translation[3]) of the
// reference's cost at `pose`, for pn correspondences: pts2d [pn,2], pts3d [pn,3], wgt2d [pn,3] (wxx,wxy,wyy), K [3,3]
void ref_pnp_cost(const double *pts2d, const double *pts3d, const double *wgt2d, const double *K, const double *pose, int pn,
                  double *residuals, double *jacobian)
{
    typedef ceres::Jet<double, 6> J;
    for (int i = 0; i < pn; ++i) {
        // same argument order as the reference's call site (uncertainty_pnp.cpp:76-80)
        ReprojectionErrorArray f(pts2d[i * 2], pts2d[i * 2 + 1], pts3d[i * 3], pts3d[i * 3 + 1], pts3d[i * 3 + 2],
                                 wgt2d[i * 3 + 0], wgt2d[i * 3 + 1], wgt2d[i * 3 + 2], K[0], K[4], K[2], K[5]);
        J p[6], r[2];
        for (int k = 0; k < 6; ++k) { p[k].a = pose[k]; p[k].v.setZero(); p[k].v[k] = 1.0; }
        f(p, r);
        for (int c = 0; c < 2; ++c) {
            residuals[i * 2 + c] = r[c].a;
            for (int k = 0; k < 6; ++k) jacobian[(i * 2 + c) * 6 + k] = r[c].v[k];
        }
    }
}

}  // extern "C"
