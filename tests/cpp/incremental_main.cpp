// incremental_main.cpp — test driver for host/adjuster/IncrementalBundleAdjuster.h: grows a WorldStructure the way SequentialActuator
// does (src/actuator/SequentialActuator.h:138-229: add an image, its LocalFrame, the points it triangulates and its observations of
// older points, then run a full BA) and adjusts after every stage, either with a fresh BundleAdjuster per call (the reference's
// pattern, :227) or with ONE IncrementalBundleAdjuster kept alive.  Both must leave the same world behind.
// usage: incremental_main <shape> <out.txt> <mode: fresh|incremental> <scenario 0|1|2> <max_iterations> <stage cuts...>
//   scenario 1: after stage 1 the world moves point 0 (an edit outside BA);  scenario 2: after stage 1 point 0 loses its last observation
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../3dreconstruction_b200/host/adjuster/BundleAdjuster.h"
#include "../../3dreconstruction_b200/host/adjuster/IncrementalBundleAdjuster.h"
#include "../../include/ba_synth.h"

int main(int argc, char** argv) {
    if (argc < 7) { std::fprintf(stderr, "usage: %s shape out mode scenario max_iterations cuts...\n", argv[0]); return 2; }
    const bool incremental = std::strcmp(argv[3], "incremental") == 0;
    const int scenario = std::atoi(argv[4]), max_it = std::atoi(argv[5]);
    std::vector<int> cuts;
    for (int a = 6; a < argc; ++a) cuts.push_back(std::atoi(argv[a]));
    ba_synth_spec spec;
    if (ba_synth_named(argv[1], &spec) != 0) return 2;
    ba_synth_problem* S = nullptr;
    if (ba_synth_generate(&spec, &S) != 0) return 3;

    auto camera = std::make_shared<Camera>(S->intr[0], S->intr[1], S->intr[2], S->intr[3]);
    std::vector<Image::Ptr> images;
    for (int c = 0; c < S->num_cameras; ++c) {
        auto img = std::make_shared<Image>("view" + std::to_string(c), camera);
        double R[9];
        b200ba::angle_axis_to_rotation(S->pose + 6 * c, R);
        img->setTcw(Sophus::SE3d(R, Eigen::Vector3d(S->pose[6 * c + 3], S->pose[6 * c + 4], S->pose[6 * c + 5])));
        images.push_back(img);
    }
    auto world = std::make_shared<WorldStructure>();
    std::vector<WorldPoint::Ptr> point_of(S->num_points);          // synthetic point id -> world point, once created
    FILE* f = std::fopen(argv[2], "w");
    if (!f) return 4;
    IncrementalBundleAdjuster inc{};
    inc.options().max_num_iterations = max_it; inc.options().num_threads = 4;
    int nc_prev = 0;
    for (size_t s = 0; s < cuts.size(); ++s) {
        const int nc = cuts[s];
        for (int c = std::max(nc_prev, 1); c < nc; ++c) world->addLocalFrame(std::make_shared<LocalFrame>(images[c - 1], images[c]));
        for (int o = 0; o < S->num_observations; ++o) {              // observations of the images added in this stage, in file order
            const int c = S->obs_cam[o], p = S->obs_pt[o];
            if (c < nc_prev || c >= nc) continue;
            if (!point_of[p]) point_of[p] = world->getPointFromIdx(world->addPoint(Eigen::Vector3d(S->pt[3 * p], S->pt[3 * p + 1], S->pt[3 * p + 2])));
            point_of[p]->observed_frames_.emplace_back(images[c], cv::Point2d(S->obs_uv[2 * o], S->obs_uv[2 * o + 1]));
        }
        if (s == 2 && scenario == 1) { auto wp = world->getPointFromIdx(0); wp->world_pos_(0) += 0.01; }
        if (s == 2 && scenario == 2) { auto wp = world->getPointFromIdx(0); wp->observed_frames_.pop_back(); }
        ba_summary sum{};
        if (incremental) { inc(world); sum = inc.summary(); }
        else { BundleAdjuster adj{}; adj.options().max_num_iterations = max_it; adj.options().num_threads = 4; adj(world); sum = adj.summary(); }
        std::fprintf(f, "stage %d %d %d %.17g %.17g\n", nc, sum.usable, sum.num_iterations, sum.initial_cost, sum.final_cost);
        nc_prev = nc;
    }
    std::fprintf(f, "calls %d %d\n", inc.appendedCalls(), inc.rebuiltCalls());
    std::fprintf(f, "intr %.17g %.17g %.17g %.17g\n", camera->get_fx(), camera->get_fy(), camera->get_cx(), camera->get_cy());
    for (int c = 0; c < nc_prev; ++c) {
        const double* R = images[c]->getTcw().rotationMatrix();
        const auto t = images[c]->getTcw().translation();
        std::fprintf(f, "pose");
        for (int k = 0; k < 9; ++k) std::fprintf(f, " %.17g", R[k]);
        std::fprintf(f, " %.17g %.17g %.17g\n", t[0], t[1], t[2]);
    }
    for (int p = 0; p < S->num_points; ++p)
        if (point_of[p]) std::fprintf(f, "pt %.17g %.17g %.17g\n", point_of[p]->world_pos_[0], point_of[p]->world_pos_[1], point_of[p]->world_pos_[2]);
    std::fclose(f);
    ba_synth_free(S);
    return 0;
}
