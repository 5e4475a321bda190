/*
 * ba_synth.h — deterministic synthetic BAL-shape problem generator (host only, no CUDA).
 *
 * The reference ships no data and no tests, so every parity and benchmark input is
 * produced here, identically for the CUDA engine and the CPU oracle.  Output is already in the
 * canonical indexing of include/ba_engine.h (observations sorted by point id).
 *
 * Points are numbered in order of their lowest-index observing camera ("birth camera"), as the
 * reference's incremental pipeline creates them: WorldStructure::addPoint hands out idx in creation
 * order (src/world/WorldStructure.h:47-52) and savePointCloudToWorld creates points when their frame
 * is added (src/actuator/SequentialActuator.h:25-72).
 */
#ifndef BA_SYNTH_H
#define BA_SYNTH_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ba_synth_spec {
    int32_t num_cameras;
    int32_t num_points;
    int32_t num_observations;   /* exact; must satisfy 2*Np <= No <= Np*min(Nc,max_track) */
    uint64_t seed;
    int32_t max_track;          /* 40 */
    double  fx, fy, cx, cy;     /* 2905.88, 2905.88, 1416, 1064 (src/main.cpp:59) */
    double  noise_px;           /* 0.5 */
    double  outlier_fraction;   /* 0.02 */
    double  outlier_px;         /* 30 */
    double  sigma_aa;           /* 0.01 rad initial-guess perturbation */
    double  sigma_t_rel;        /* 0.01 of scene extent */
    double  sigma_pt_rel;       /* 0.02 of scene extent */
    int32_t round_obs_to_float; /* 1: observations pass through float like cv::KeyPoint::pt (SequentialActuator.h:41,66-67) */
    int32_t first_pose_zero;    /* 1: scene expressed in camera 0's frame, pose 0 exactly zero (reference flow, F8b) */
    int32_t fixed_camera;       /* index of the constant pose (reference: 1, the first frame's image2, F8a); -1 none */
    int32_t arc;                /* 1: cameras on a 120 degree arc (sequential-30); 0: closed ring */
} ba_synth_spec;

typedef struct ba_synth_problem {
    int32_t num_cameras, num_groups, num_points, num_observations;
    double*  pose;       /* [Nc][6] initial guess */
    double*  intr;       /* [1][4] */
    int32_t* cam_group;  /* [Nc] all 0 */
    uint8_t* cam_fixed;  /* [Nc] */
    double*  pt;         /* [Np][3] initial guess */
    int32_t* obs_cam;    /* [No] */
    int32_t* obs_pt;     /* [No] sorted */
    double*  obs_uv;     /* [No][2] */
    double*  pose_true;  /* [Nc][6] */
    double*  pt_true;    /* [Np][3] */
    double   extent;
} ba_synth_problem;

int  ba_synth_spec_default(ba_synth_spec* s);
/* Named shapes of BASELINE.json: "ladybug" 49/7776/31843, "trafalgar" 257/65132/225911,
 * "venice" 1778/993923/5001946, "final" 13682/4456117/28987644, "sequential30" 30/~7k/~25k, "tiny" 6/40/140. */
int  ba_synth_named(const char* name, ba_synth_spec* s);
int  ba_synth_generate(const ba_synth_spec* s, ba_synth_problem** out);
void ba_synth_free(ba_synth_problem* p);

/* "Bundle Adjustment in the Large" text files (the format of the reference's SnavelyReprojectionError model,
 * src/adjuster/SnavelyReprojectionError.h:16-55): 9 parameters per camera (angle-axis, t, f, k1, k2), 3 per point.
 * ba_bal_read returns a problem with one intrinsics group per camera (f, k1, k2, 0), observations stably sorted by point,
 * nothing fixed; free it with ba_synth_free.  Solve with ba_options.model = BA_MODEL_SNAVELY. */
int  ba_bal_read(const char* path, ba_synth_problem** out);
int  ba_bal_write(const char* path, int32_t num_cameras, int32_t num_points, int32_t num_observations, const double* pose, const double* intr,
                  const int32_t* cam_group, const double* pt, const int32_t* obs_cam, const int32_t* obs_pt, const double* obs_uv);
/* ASCII PLY of points (white) and camera centres (green), the reference's exportToPly / writeToPLYFile
 * (src/helpers/SfMPlyHelper.hpp:62-116, src/world/WorldStructure.h:83-98). */
int  ba_write_ply(const char* path, int32_t num_points, const double* pt, int32_t num_cameras, const double* pose);

#ifdef __cplusplus
}
#endif
#endif
