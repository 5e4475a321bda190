// adapter_main.cpp — test driver for host/adjuster/BundleAdjuster.h: builds a WorldStructure from a synthetic scene the way
// SequentialActuator would hand it to BA (src/actuator/SequentialActuator.h:25-72,85-136: frames (image[i-1], image[i]),
// first image at identity and never loaded, points in creation order with their observed_frames_), runs
//     BundleAdjuster adj; adj(world);
// and dumps the world afterwards.  Linked by tests/test_adapter.py against libba_engine.so (GPU box) or, on the CPU box,
// against the oracle library that implements the same header (checker use only).
// usage: adapter_main <shape> <out.txt> [euler_bug 0|1] [intrinsics_free 0|1] [max_iterations]
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../3dreconstruction_b200/host/adjuster/BundleAdjuster.h"
#include "../../include/ba_synth.h"

int main(int argc, char** argv) {
    if (argc < 3) { std::fprintf(stderr, "usage: %s shape out [euler_bug] [intrinsics_free] [max_iterations]\n", argv[0]); return 2; }
    const bool euler_bug = argc > 3 && std::atoi(argv[3]) != 0;
    const int fi = argc > 4 ? std::atoi(argv[4]) : 1;
    const int max_it = argc > 5 ? std::atoi(argv[5]) : 50;
    ba_synth_spec spec;
    if (ba_synth_named(argv[1], &spec) != 0) { std::fprintf(stderr, "unknown shape %s\n", argv[1]); return 2; }
    ba_synth_problem* S = nullptr;
    if (ba_synth_generate(&spec, &S) != 0) return 3;

    auto camera = std::make_shared<Camera>(S->intr[0], S->intr[1], S->intr[2], S->intr[3]);   // ONE shared Camera (main.cpp:91)
    std::vector<Image::Ptr> images;
    for (int c = 0; c < S->num_cameras; ++c) {
        auto img = std::make_shared<Image>("view" + std::to_string(c), camera);
        double R[9];
        b200ba::angle_axis_to_rotation(S->pose + 6 * c, R);
        img->setTcw(Sophus::SE3d(R, Eigen::Vector3d(S->pose[6 * c + 3], S->pose[6 * c + 4], S->pose[6 * c + 5])));
        images.push_back(img);
    }
    auto world = std::make_shared<WorldStructure>();
    for (int c = 1; c < S->num_cameras; ++c) world->addLocalFrame(std::make_shared<LocalFrame>(images[c - 1], images[c]));
    int o = 0;
    for (int p = 0; p < S->num_points; ++p) {
        const auto idx = world->addPoint(Eigen::Vector3d(S->pt[3 * p], S->pt[3 * p + 1], S->pt[3 * p + 2]));
        auto wp = world->getPointFromIdx(idx);
        for (; o < S->num_observations && S->obs_pt[o] == p; ++o)
            wp->observed_frames_.emplace_back(images[S->obs_cam[o]], cv::Point2d(S->obs_uv[2 * o], S->obs_uv[2 * o + 1]));
    }

    Image::write_back = euler_bug ? Image::WriteBack::ReferenceEulerBug : Image::WriteBack::CorrectAngleAxis;
    BundleAdjuster adj{};
    adj.options().intrinsics_free = fi;
    adj.options().max_num_iterations = max_it;
    adj.options().num_threads = 4;   // read by the oracle only
    adj(world);

    FILE* f = std::fopen(argv[2], "w");
    if (!f) return 4;
    const ba_summary& s = adj.summary();
    std::fprintf(f, "summary %d %d %d %.17g %.17g\n", s.termination, s.usable, s.num_iterations, s.initial_cost, s.final_cost);
    std::fprintf(f, "intr %.17g %.17g %.17g %.17g\n", camera->get_fx(), camera->get_fy(), camera->get_cx(), camera->get_cy());
    for (auto& img : images) {
        const double* R = img->getTcw().rotationMatrix();
        const auto t = img->getTcw().translation();
        std::fprintf(f, "pose");
        for (int k = 0; k < 9; ++k) std::fprintf(f, " %.17g", R[k]);
        std::fprintf(f, " %.17g %.17g %.17g\n", t[0], t[1], t[2]);
    }
    for (int p = 0; p < S->num_points; ++p) {
        const auto x = world->getPointFromIdx(size_t(p))->world_pos_;
        std::fprintf(f, "pt %.17g %.17g %.17g\n", x[0], x[1], x[2]);
    }
    for (auto& r : adj.trajectory()) std::fprintf(f, "iter %d %d %.17g\n", r.iteration, r.step_is_successful, r.cost);
    std::fclose(f);
    ba_synth_free(S);
    return 0;
}
