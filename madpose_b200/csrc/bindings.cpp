// pybind11 module `_madpose` (re-exported by madpose_b200/__init__.py): the drop-in Python surface of
// /root/reference/src/bindings.cpp:22-175 -- same class, field and function names, same argument order,
// same returned (model, statistics) tuples -- in front of the C ABI of include/madpose_b200.h.
// Host C++ only (no Eigen, no CUDA headers): numpy in, numpy out, the GIL released around the device call.
// Everything the reference binds is bound here with a device body, including the scale-only estimator
// HybridEstimatePoseAndScale / PoseAndScale (bindings.cpp:111-118,167-168) and the seven stand-alone solver
// functions (bindings.cpp:156-166).
#include <pybind11/iostream.h>
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/madpose_b200.h"

namespace py = pybind11;
using namespace pybind11::literals;
typedef py::array_t<double, py::array::c_style | py::array::forcecast> darray;

namespace {

// ---- option / statistics records (names: ext/RansacLib/RansacLib/hybrid_ransac.h:47-104, src/hybrid_ransac.h:17-25)
struct HybridLORansacOptions {
    uint32_t min_num_iterations = 100, max_num_iterations = 10000, max_num_iterations_per_solver = 10000;
    double success_probability = 0.9999;
    std::vector<double> squared_inlier_thresholds, data_type_weights;
    unsigned int random_seed = 0;
    int num_lo_steps = 10;
    double threshold_multiplier = 1.4142135623730951;
    int num_lsq_iterations = 4, min_sample_multiplicator = 7, non_min_sample_multiplier = 3;
    uint32_t lo_starting_iterations = 50;
    bool final_least_squares = false, use_ours = false, use_4p4d = false;
};
struct EstimatorConfig { // src/estimator_config.h:5-33
    int solver_type = 0, score_type = 0, LO_type = 0;
    bool min_depth_constraint = true, use_shift = true;
    double ceres_function_tolerance = 1e-6, ceres_gradient_tolerance = 1e-8, ceres_parameter_tolerance = 1e-6;
    double ceres_max_num_iterations = 25;
    bool ceres_use_nonmonotonic_steps = true;
    int ceres_num_threads = 1;
    EstimatorConfig() {}
    EstimatorConfig(int solver, int score, int lo) : solver_type(solver), score_type(score), LO_type(lo) {}
};
struct HybridRansacStatistics {
    uint32_t num_iterations_total = 0;
    std::vector<uint32_t> num_iterations_per_solver;
    int best_num_inliers = 0, best_solver_type = -1;
    double best_model_score = 1.7976931348623157e308;
    std::vector<double> inlier_ratios;
    std::vector<std::vector<int>> inlier_indices;
    int number_lo_iterations = 0;
};
struct RansacStats { // ransac_lib::RansacStatistics: bound by the reference for API completeness only
    uint32_t num_iterations = 0;
    int best_num_inliers = 0;
    double best_model_score = 1.7976931348623157e308, inlier_ratio = 0;
    std::vector<int> inlier_indices;
    int number_lo_iterations = 0;
};
struct RansacOptions {
    uint32_t min_num_iterations_ = 100, max_num_iterations_ = 10000;
    double success_probability_ = 0.9999, squared_inlier_threshold_ = 1.0;
    unsigned int random_seed_ = 0;
};
struct LORansacOptions {
    uint32_t min_num_iterations = 100, max_num_iterations = 10000;
    double success_probability = 0.9999, squared_inlier_threshold = 1.0;
    unsigned int random_seed = 0;
    int num_lo_steps = 10;
    double threshold_multiplier = 1.4142135623730951;
    int num_lsq_iterations = 4, min_sample_multiplicator = 7, non_min_sample_multiplier = 3;
    uint32_t lo_starting_iterations = 50;
    bool final_least_squares = false;
};

// ---- model records (src/pose.h:7-56): 3x4 pose [R | t] row-major
struct PoseScaleOffset {
    double pose[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    double scale = 1, offset0 = 0, offset1 = 0;
    void set(const darray &p) {
        if (p.size() != 12) throw std::invalid_argument("pose must be 3x4");
        std::memcpy(pose, p.data(), sizeof(pose));
    }
    void set(const darray &R, const darray &t) {
        if (R.size() != 9 || t.size() != 3) throw std::invalid_argument("R must be 3x3 and t a 3-vector");
        for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) pose[4 * i + j] = R.data()[3 * i + j]; pose[4 * i + 3] = t.data()[i]; }
    }
    py::array get_pose() const { darray a({3, 4}); std::memcpy(a.mutable_data(), pose, sizeof(pose)); return a; }
    py::array R() const { darray a({3, 3}); for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) a.mutable_data()[3 * i + j] = pose[4 * i + j]; return a; }
    py::array t() const { darray a(3); for (int i = 0; i < 3; i++) a.mutable_data()[i] = pose[4 * i + 3]; return a; }
    void from_model(const mp_model &m) {
        for (int i = 0; i < 3; i++) { for (int j = 0; j < 3; j++) pose[4 * i + j] = m.R[3 * i + j]; pose[4 * i + 3] = m.t[i]; }
        scale = m.scale; offset0 = m.offset0; offset1 = m.offset1;
    }
};
struct PoseScaleOffsetSharedFocal : PoseScaleOffset { double focal = 1; };
struct PoseAndScale { // pose.h:7-17: the base record, no offsets
    PoseScaleOffset b;
    double scale = 1;
};
struct PoseScaleOffsetTwoFocal : PoseScaleOffset { double focal0 = 1, focal1 = 1; };

// ---- one device context per process (created on first use; MADPOSE_B200_DEVICE selects the GPU)
mp_ctx *g_ctx = nullptr;
std::mutex g_mu;
mp_ctx *context() {
    if (!g_ctx) {
        const char *dev = std::getenv("MADPOSE_B200_DEVICE");
        if (mp_create(dev ? std::atoi(dev) : 0, &g_ctx) != MP_OK)
            throw std::runtime_error(std::string("madpose_b200: ") + mp_last_error());
    }
    return g_ctx;
}

mp_options to_c(const HybridLORansacOptions &o) {
    if (o.squared_inlier_thresholds.size() != 2 || o.data_type_weights.size() != 2)
        throw std::invalid_argument("squared_inlier_thresholds and data_type_weights must have 2 entries "
                                    "(the reference indexes [1] unconditionally, hybrid_pose_estimator.cpp:16-17)");
    mp_options c;
    std::memset(&c, 0, sizeof(c));
    c.min_num_iterations = o.min_num_iterations; c.max_num_iterations = o.max_num_iterations;
    c.max_num_iterations_per_solver = o.max_num_iterations_per_solver; c.random_seed = o.random_seed;
    c.success_probability = o.success_probability;
    for (int k = 0; k < 2; k++) { c.squared_inlier_thresholds[k] = o.squared_inlier_thresholds[k]; c.data_type_weights[k] = o.data_type_weights[k]; }
    c.threshold_multiplier = o.threshold_multiplier; c.num_lo_steps = o.num_lo_steps; c.num_lsq_iterations = o.num_lsq_iterations;
    c.min_sample_multiplicator = o.min_sample_multiplicator; c.non_min_sample_multiplier = o.non_min_sample_multiplier;
    c.lo_starting_iterations = o.lo_starting_iterations; c.final_least_squares = o.final_least_squares;
    c.use_ours = o.use_ours; c.use_4p4d = o.use_4p4d;
    return c;
}
mp_config to_c(const EstimatorConfig &e) {
    mp_config c;
    std::memset(&c, 0, sizeof(c));
    c.solver_type = e.solver_type; c.score_type = e.score_type; c.lo_type = e.LO_type;
    c.min_depth_constraint = e.min_depth_constraint; c.use_shift = e.use_shift;
    c.ceres_use_nonmonotonic_steps = e.ceres_use_nonmonotonic_steps; c.ceres_num_threads = e.ceres_num_threads;
    c.ceres_function_tolerance = e.ceres_function_tolerance; c.ceres_gradient_tolerance = e.ceres_gradient_tolerance;
    c.ceres_parameter_tolerance = e.ceres_parameter_tolerance; c.ceres_max_num_iterations = e.ceres_max_num_iterations;
    return c;
}

// ---- the shared body of the three estimators: a batch of pairs through mp_estimate_batch
struct PairIn { darray x0, x1, d0, d1, min_depth, A, B; }; // A,B = K0,K1 (3x3) or pp0,pp1 (2)

void run_batch(int kind, std::vector<PairIn> &in, const HybridLORansacOptions &options, const EstimatorConfig &cfg,
               std::vector<mp_model> *models, std::vector<HybridRansacStatistics> *stats, const std::vector<int> &devices = {}) {
    const int P = (int)in.size();
    std::vector<mp_pair> pairs(P);
    size_t total = 0;
    for (int p = 0; p < P; p++) {
        PairIn &q = in[p];
        if (q.x0.ndim() != 2 || q.x0.shape(1) != 2 || q.x1.ndim() != 2 || q.x1.shape(1) != 2)
            throw std::invalid_argument("x0 and x1 must be (N, 2)");
        const py::ssize_t n = q.x0.shape(0);
        if (q.x1.shape(0) != n || q.d0.size() != n || q.d1.size() != n)
            throw std::invalid_argument("x0, x1, depth0, depth1 must have the same length"); // UB in the reference
        if (kind != MP_KIND_SCALE_ONLY && q.min_depth.size() != 2) throw std::invalid_argument("min_depth must have 2 entries");
        mp_pair &c = pairs[p];
        std::memset(&c, 0, sizeof(c));
        c.n = (int)n; c.x0 = q.x0.data(); c.x1 = q.x1.data(); c.d0 = q.d0.data(); c.d1 = q.d1.data();
        if (kind != MP_KIND_SCALE_ONLY) { c.min_depth[0] = q.min_depth.data()[0]; c.min_depth[1] = q.min_depth.data()[1]; }
        if (kind == MP_KIND_CALIBRATED || kind == MP_KIND_SCALE_ONLY) {
            if (q.A.size() != 9 || q.B.size() != 9) throw std::invalid_argument("K0 and K1 must be 3x3");
            std::memcpy(c.K0, q.A.data(), 72); std::memcpy(c.K1, q.B.data(), 72);
        } else {
            if (q.A.size() != 2 || q.B.size() != 2) throw std::invalid_argument("pp0 and pp1 must be 2-vectors");
            c.pp0[0] = q.A.data()[0]; c.pp0[1] = q.A.data()[1]; c.pp1[0] = q.B.data()[0]; c.pp1[1] = q.B.data()[1];
        }
        total += (size_t)n;
    }
    const mp_options o = to_c(options);
    const mp_config e = to_c(cfg);
    models->resize(P);
    std::vector<mp_stats> st(P);
    std::vector<int32_t> inl(3 * total + 1);
    int rc;
    std::string err;
    {
        py::gil_scoped_release nogil; // the reference holds the GIL for the whole call; nothing here touches Python
        if (devices.size() > 1) { // the batch is sharded over several GPUs inside the library (no inter-GPU traffic)
            rc = mp_estimate_batch_multi((int)devices.size(), devices.data(), kind, P, pairs.data(), &o, &e, models->data(), st.data(), inl.data());
        } else {
            std::lock_guard<std::mutex> lock(g_mu);
            rc = mp_estimate_batch(context(), kind, P, pairs.data(), &o, &e, models->data(), st.data(), inl.data());
        }
        if (rc != MP_OK) err = mp_last_error();
    }
    if (rc == MP_ERR_ARG) throw std::invalid_argument("madpose_b200: " + err);
    if (rc != MP_OK) throw std::runtime_error("madpose_b200: " + err);
    stats->resize(P);
    size_t off = 0;
    for (int p = 0; p < P; p++) {
        HybridRansacStatistics &h = (*stats)[p];
        const mp_stats &s = st[p];
        h.num_iterations_total = s.num_iterations_total;
        h.num_iterations_per_solver = {s.num_iterations_per_solver[0], s.num_iterations_per_solver[1]};
        h.best_num_inliers = s.best_num_inliers; h.best_solver_type = s.best_solver_type;
        h.best_model_score = s.best_model_score; h.number_lo_iterations = s.number_lo_iterations;
        h.inlier_ratios = {s.inlier_ratios[0], s.inlier_ratios[1], s.inlier_ratios[2]};
        h.inlier_indices.assign(3, std::vector<int>());
        for (int t = 0; t < 3; t++)
            h.inlier_indices[t].assign(inl.begin() + off + (size_t)t * pairs[p].n, inl.begin() + off + (size_t)t * pairs[p].n + s.num_inliers[t]);
        off += 3 * (size_t)pairs[p].n;
    }
}

template <class M> M to_py(int kind, const mp_model &m);
template <> PoseScaleOffset to_py<PoseScaleOffset>(int, const mp_model &m) { PoseScaleOffset r; r.from_model(m); return r; }
template <> PoseScaleOffsetSharedFocal to_py<PoseScaleOffsetSharedFocal>(int, const mp_model &m) {
    PoseScaleOffsetSharedFocal r; r.from_model(m); r.focal = m.focal0; return r;
}
template <> PoseScaleOffsetTwoFocal to_py<PoseScaleOffsetTwoFocal>(int, const mp_model &m) {
    PoseScaleOffsetTwoFocal r; r.from_model(m); r.focal0 = m.focal0; r.focal1 = m.focal1; return r;
}

template <> PoseAndScale to_py<PoseAndScale>(int, const mp_model &m) { PoseAndScale r; r.b.from_model(m); r.scale = m.scale; return r; }

// ---- the stand-alone solver functions (bindings.cpp:156-166) through mp_solve_affine / mp_estimate_scale_and_pose
void check_sample(int kind, const darray &x, const darray &y, const darray &dx, const darray &dy) {
    const py::ssize_t np_ = kind == MP_KIND_CALIBRATED ? 3 : 4;
    if (x.ndim() != 2 || x.shape(0) != 3 || x.shape(1) != np_ || y.ndim() != 2 || y.shape(0) != 3 || y.shape(1) != np_)
        throw std::invalid_argument(kind == MP_KIND_CALIBRATED ? "x_homo and y_homo must be 3x3 (columns = homogeneous points)"
                                                                : "x_homo and y_homo must be 3x4 (columns = homogeneous points)");
    if (dx.size() != np_ || dy.size() != np_) throw std::invalid_argument("depth_x and depth_y must have one entry per point");
}
void solve_affine(int kind, const darray &x, const darray &y, const darray &dx, const darray &dy, double *sols, int *ns,
                  mp_model *models, int *nm) {
    check_sample(kind, x, y, dx, dy);
    int rc;
    std::string err;
    {
        py::gil_scoped_release nogil;
        std::lock_guard<std::mutex> lock(g_mu);
        rc = mp_solve_affine(context(), kind, 1, x.data(), y.data(), dx.data(), dy.data(), sols, ns, models, nm);
        if (rc != MP_OK) err = mp_last_error();
    }
    if (rc == MP_ERR_ARG) throw std::invalid_argument("madpose_b200: " + err);
    if (rc != MP_OK) throw std::runtime_error("madpose_b200: " + err);
}
// solve_scale_and_shift* -> list of (1, b1, a2, b2 a2[, f | f1, f2]) vectors (solver.cpp:117-121, 285-289, 472-477)
template <int KIND> std::vector<py::array> solve_raw(darray x, darray y, darray dx, darray dy) {
    double sols[48];
    int ns = 0;
    solve_affine(KIND, x, y, dx, dy, sols, &ns, nullptr, nullptr);
    const int w = KIND == MP_KIND_CALIBRATED ? 4 : (KIND == MP_KIND_SHARED_FOCAL ? 5 : 6);
    std::vector<py::array> out;
    for (int q = 0; q < ns; q++) {
        darray a(w);
        for (int k = 0; k < w; k++) a.mutable_data()[k] = sols[6 * q + k];
        out.push_back(a);
    }
    return out;
}
// solve_scale_shift_pose*_wrapper -> list of pose records (scale = a2, offset0 = b1, offset1 = b2 a2)
template <class M, int KIND> std::vector<M> solve_pose(darray x, darray y, darray dx, darray dy) {
    mp_model models[8];
    int nm = 0;
    solve_affine(KIND, x, y, dx, dy, nullptr, nullptr, models, &nm);
    std::vector<M> out;
    for (int q = 0; q < nm; q++) out.push_back(to_py<M>(KIND, models[q]));
    return out;
}
PoseAndScale scale_and_pose(darray X, darray Y, darray W) {
    if (X.ndim() != 2 || X.shape(0) != 3 || Y.ndim() != 2 || Y.shape(0) != 3 || X.shape(1) != Y.shape(1) || W.size() != X.shape(1))
        throw std::invalid_argument("X and Y must be 3xN and W must have N entries");
    mp_model m;
    int rc;
    std::string err;
    {
        py::gil_scoped_release nogil;
        std::lock_guard<std::mutex> lock(g_mu);
        rc = mp_estimate_scale_and_pose(context(), (int)X.shape(1), X.data(), Y.data(), W.data(), &m);
        if (rc != MP_OK) err = mp_last_error();
    }
    if (rc == MP_ERR_ARG) throw std::invalid_argument("madpose_b200: " + err);
    if (rc != MP_OK) throw std::runtime_error("madpose_b200: " + err);
    return to_py<PoseAndScale>(0, m);
}

template <class M, int KIND>
std::pair<M, HybridRansacStatistics> estimate_one(darray x0, darray x1, darray d0, darray d1, darray min_depth, darray A,
                                                  darray B, const HybridLORansacOptions &options, const EstimatorConfig &cfg) {
    std::vector<PairIn> in(1);
    in[0] = PairIn{x0, x1, d0, d1, min_depth, A, B};
    std::vector<mp_model> models;
    std::vector<HybridRansacStatistics> stats;
    run_batch(KIND, in, options, cfg, &models, &stats);
    return {to_py<M>(KIND, models[0]), stats[0]};
}
// batch form (superset API): lists of per-pair arrays -> list of (model, statistics)
template <class M, int KIND>
std::vector<std::pair<M, HybridRansacStatistics>> estimate_many(std::vector<darray> x0, std::vector<darray> x1,
                                                                std::vector<darray> d0, std::vector<darray> d1,
                                                                std::vector<darray> min_depth, std::vector<darray> A,
                                                                std::vector<darray> B, const HybridLORansacOptions &options,
                                                                const EstimatorConfig &cfg, const std::vector<int> &devices = {}) {
    const size_t P = x0.size();
    if (x1.size() != P || d0.size() != P || d1.size() != P || min_depth.size() != P || A.size() != P || B.size() != P)
        throw std::invalid_argument("all per-pair lists must have the same length");
    std::vector<PairIn> in(P);
    for (size_t p = 0; p < P; p++) in[p] = PairIn{x0[p], x1[p], d0[p], d1[p], min_depth[p], A[p], B[p]};
    std::vector<mp_model> models;
    std::vector<HybridRansacStatistics> stats;
    run_batch(KIND, in, options, cfg, &models, &stats, devices);
    std::vector<std::pair<M, HybridRansacStatistics>> out;
    for (size_t p = 0; p < P; p++) out.push_back({to_py<M>(KIND, models[p]), stats[p]});
    return out;
}

// HybridEstimatePoseAndScale (bindings.cpp:167-168): no min_depth argument
std::pair<PoseAndScale, HybridRansacStatistics> estimate_scale_only(darray x0, darray x1, darray d0, darray d1, darray K0, darray K1,
                                                                    const HybridLORansacOptions &options, const EstimatorConfig &cfg) {
    return estimate_one<PoseAndScale, MP_KIND_SCALE_ONLY>(x0, x1, d0, d1, darray(), K0, K1, options, cfg);
}
std::vector<std::pair<PoseAndScale, HybridRansacStatistics>> estimate_scale_only_many(std::vector<darray> x0, std::vector<darray> x1,
                                                                                      std::vector<darray> d0, std::vector<darray> d1,
                                                                                      std::vector<darray> K0, std::vector<darray> K1,
                                                                                      const HybridLORansacOptions &options,
                                                                                      const EstimatorConfig &cfg, const std::vector<int> &devices = {}) {
    std::vector<darray> md(x0.size());
    return estimate_many<PoseAndScale, MP_KIND_SCALE_ONLY>(x0, x1, d0, d1, md, K0, K1, options, cfg, devices);
}

template <class M> void bind_pose_common(py::class_<M> &c) {
    c.def_property("pose", [](const M &m) { return m.get_pose(); }, [](M &m, const darray &p) { m.set(p); })
        .def_readwrite("scale", &M::scale)
        .def_readwrite("offset0", &M::offset0)
        .def_readwrite("offset1", &M::offset1)
        .def("R", &M::R)
        .def("t", &M::t);
}

} // namespace

PYBIND11_MODULE(_madpose, m) {
    m.doc() = "B200-native hybrid LO-RANSAC path of MADPose (relative pose through affine correction of monocular "
              "depth priors) behind the reference's Python API.";
    py::add_ostream_redirect(m, "ostream_redirect");

    py::class_<RansacStats>(m, "RansacStats")
        .def(py::init<>())
        .def_readwrite("num_iterations", &RansacStats::num_iterations)
        .def_readwrite("best_num_inliers", &RansacStats::best_num_inliers)
        .def_readwrite("best_model_score", &RansacStats::best_model_score)
        .def_readwrite("inlier_ratio", &RansacStats::inlier_ratio)
        .def_readwrite("inlier_indices", &RansacStats::inlier_indices)
        .def_readwrite("number_lo_iterations", &RansacStats::number_lo_iterations);
    py::class_<RansacOptions>(m, "RansacOptions")
        .def(py::init<>())
        .def_readwrite("min_num_iterations_", &RansacOptions::min_num_iterations_)
        .def_readwrite("max_num_iterations_", &RansacOptions::max_num_iterations_)
        .def_readwrite("success_probability_", &RansacOptions::success_probability_)
        .def_readwrite("squared_inlier_threshold_", &RansacOptions::squared_inlier_threshold_)
        .def_readwrite("random_seed_", &RansacOptions::random_seed_);
    py::class_<LORansacOptions>(m, "LORansacOptions")
        .def(py::init<>())
        .def_readwrite("min_num_iterations", &LORansacOptions::min_num_iterations)
        .def_readwrite("max_num_iterations", &LORansacOptions::max_num_iterations)
        .def_readwrite("success_probability", &LORansacOptions::success_probability)
        .def_readwrite("squared_inlier_threshold", &LORansacOptions::squared_inlier_threshold)
        .def_readwrite("random_seed", &LORansacOptions::random_seed)
        .def_readwrite("num_lo_steps", &LORansacOptions::num_lo_steps)
        .def_readwrite("threshold_multiplier", &LORansacOptions::threshold_multiplier)
        .def_readwrite("num_lsq_iterations", &LORansacOptions::num_lsq_iterations)
        .def_readwrite("min_sample_multiplicator", &LORansacOptions::min_sample_multiplicator)
        .def_readwrite("non_min_sample_multiplier", &LORansacOptions::non_min_sample_multiplier)
        .def_readwrite("lo_starting_iterations", &LORansacOptions::lo_starting_iterations)
        .def_readwrite("final_least_squares", &LORansacOptions::final_least_squares);
    py::class_<HybridRansacStatistics>(m, "HybridRansacStatistics")
        .def(py::init<>())
        .def_readwrite("num_iterations_total", &HybridRansacStatistics::num_iterations_total)
        .def_readwrite("num_iterations_per_solver", &HybridRansacStatistics::num_iterations_per_solver)
        .def_readwrite("best_num_inliers", &HybridRansacStatistics::best_num_inliers)
        .def_readwrite("best_solver_type", &HybridRansacStatistics::best_solver_type)
        .def_readwrite("best_model_score", &HybridRansacStatistics::best_model_score)
        .def_readwrite("inlier_ratios", &HybridRansacStatistics::inlier_ratios)
        .def_readwrite("inlier_indices", &HybridRansacStatistics::inlier_indices)
        .def_readwrite("number_lo_iterations", &HybridRansacStatistics::number_lo_iterations);
    py::class_<HybridLORansacOptions>(m, "HybridLORansacOptions")
        .def(py::init<>())
        .def_readwrite("min_num_iterations", &HybridLORansacOptions::min_num_iterations)
        .def_readwrite("max_num_iterations", &HybridLORansacOptions::max_num_iterations)
        .def_readwrite("max_num_iterations_per_solver", &HybridLORansacOptions::max_num_iterations_per_solver)
        .def_readwrite("success_probability", &HybridLORansacOptions::success_probability)
        .def_readwrite("squared_inlier_thresholds", &HybridLORansacOptions::squared_inlier_thresholds)
        .def_readwrite("data_type_weights", &HybridLORansacOptions::data_type_weights)
        .def_readwrite("random_seed", &HybridLORansacOptions::random_seed)
        .def_readwrite("num_lo_steps", &HybridLORansacOptions::num_lo_steps)
        .def_readwrite("threshold_multiplier", &HybridLORansacOptions::threshold_multiplier)
        .def_readwrite("num_lsq_iterations", &HybridLORansacOptions::num_lsq_iterations)
        .def_readwrite("min_sample_multiplicator", &HybridLORansacOptions::min_sample_multiplicator)
        .def_readwrite("non_min_sample_multiplier", &HybridLORansacOptions::non_min_sample_multiplier)
        .def_readwrite("lo_starting_iterations", &HybridLORansacOptions::lo_starting_iterations)
        .def_readwrite("final_least_squares", &HybridLORansacOptions::final_least_squares)
        .def_readwrite("use_ours", &HybridLORansacOptions::use_ours)
        .def_readwrite("use_4p4d", &HybridLORansacOptions::use_4p4d);
    py::class_<EstimatorConfig>(m, "EstimatorConfig")
        .def(py::init<>())
        .def(py::init<int, int, int>(), "solver"_a = 0, "score"_a = 0, "LO"_a = 0)
        .def_readwrite("min_depth_constraint", &EstimatorConfig::min_depth_constraint)
        .def_readwrite("use_shift", &EstimatorConfig::use_shift)
        .def_readwrite("ceres_function_tolerance", &EstimatorConfig::ceres_function_tolerance)
        .def_readwrite("ceres_gradient_tolerance", &EstimatorConfig::ceres_gradient_tolerance)
        .def_readwrite("ceres_parameter_tolerance", &EstimatorConfig::ceres_parameter_tolerance)
        .def_readwrite("ceres_max_num_iterations", &EstimatorConfig::ceres_max_num_iterations)
        .def_readwrite("ceres_use_nonmonotonic_steps", &EstimatorConfig::ceres_use_nonmonotonic_steps)
        .def_readwrite("ceres_num_threads", &EstimatorConfig::ceres_num_threads);

    py::class_<PoseAndScale>(m, "PoseAndScale")
        .def(py::init<>())
        .def(py::init([](const darray &p, double s) { PoseAndScale r; r.b.set(p); r.scale = s; return r; }))
        .def(py::init([](const darray &R, const darray &t, double s) { PoseAndScale r; r.b.set(R, t); r.scale = s; return r; }))
        .def_property("pose", [](const PoseAndScale &m) { return m.b.get_pose(); }, [](PoseAndScale &m, const darray &p) { m.b.set(p); })
        .def_readwrite("scale", &PoseAndScale::scale)
        .def("R", [](const PoseAndScale &m) { return m.b.R(); })
        .def("t", [](const PoseAndScale &m) { return m.b.t(); });
    py::class_<PoseScaleOffset> pso(m, "PoseScaleOffset");
    pso.def(py::init<>())
        .def(py::init([](const darray &p, double s, double o0, double o1) { PoseScaleOffset r; r.set(p); r.scale = s; r.offset0 = o0; r.offset1 = o1; return r; }))
        .def(py::init([](const darray &R, const darray &t, double s, double o0, double o1) { PoseScaleOffset r; r.set(R, t); r.scale = s; r.offset0 = o0; r.offset1 = o1; return r; }));
    bind_pose_common(pso);
    py::class_<PoseScaleOffsetSharedFocal> psf(m, "PoseScaleOffsetSharedFocal");
    psf.def(py::init<>())
        .def(py::init([](const darray &p, double s, double o0, double o1, double f) { PoseScaleOffsetSharedFocal r; r.set(p); r.scale = s; r.offset0 = o0; r.offset1 = o1; r.focal = f; return r; }))
        .def(py::init([](const darray &R, const darray &t, double s, double o0, double o1, double f) { PoseScaleOffsetSharedFocal r; r.set(R, t); r.scale = s; r.offset0 = o0; r.offset1 = o1; r.focal = f; return r; }))
        .def_readwrite("focal", &PoseScaleOffsetSharedFocal::focal);
    bind_pose_common(psf);
    py::class_<PoseScaleOffsetTwoFocal> ptf(m, "PoseScaleOffsetTwoFocal");
    ptf.def(py::init<>())
        .def(py::init([](const darray &p, double s, double o0, double o1, double f0, double f1) { PoseScaleOffsetTwoFocal r; r.set(p); r.scale = s; r.offset0 = o0; r.offset1 = o1; r.focal0 = f0; r.focal1 = f1; return r; }))
        .def(py::init([](const darray &R, const darray &t, double s, double o0, double o1, double f0, double f1) { PoseScaleOffsetTwoFocal r; r.set(R, t); r.scale = s; r.offset0 = o0; r.offset1 = o1; r.focal0 = f0; r.focal1 = f1; return r; }))
        .def_readwrite("focal0", &PoseScaleOffsetTwoFocal::focal0)
        .def_readwrite("focal1", &PoseScaleOffsetTwoFocal::focal1);
    bind_pose_common(ptf);

    m.def("HybridEstimatePoseScaleOffset", &estimate_one<PoseScaleOffset, MP_KIND_CALIBRATED>, "x0"_a, "x1"_a, "depth0"_a,
          "depth1"_a, "min_depth"_a, "K0"_a, "K1"_a, "options"_a, "est_config"_a = EstimatorConfig());
    m.def("HybridEstimatePoseScaleOffsetSharedFocal", &estimate_one<PoseScaleOffsetSharedFocal, MP_KIND_SHARED_FOCAL>, "x0"_a,
          "x1"_a, "depth0"_a, "depth1"_a, "min_depth"_a, "pp0"_a, "pp1"_a, "options"_a, "est_config"_a = EstimatorConfig());
    m.def("HybridEstimatePoseScaleOffsetTwoFocal", &estimate_one<PoseScaleOffsetTwoFocal, MP_KIND_TWO_FOCAL>, "x0"_a, "x1"_a,
          "depth0"_a, "depth1"_a, "min_depth"_a, "pp0"_a, "pp1"_a, "options"_a, "est_config"_a = EstimatorConfig());
    // batch superset (one call, many pairs: what the GPU is for); arguments are lists of the per-pair arrays
    m.def("HybridEstimatePoseScaleOffsetBatch", &estimate_many<PoseScaleOffset, MP_KIND_CALIBRATED>, "x0"_a, "x1"_a, "depth0"_a,
          "depth1"_a, "min_depth"_a, "K0"_a, "K1"_a, "options"_a, "est_config"_a = EstimatorConfig(), "devices"_a = std::vector<int>());
    m.def("HybridEstimatePoseScaleOffsetSharedFocalBatch", &estimate_many<PoseScaleOffsetSharedFocal, MP_KIND_SHARED_FOCAL>,
          "x0"_a, "x1"_a, "depth0"_a, "depth1"_a, "min_depth"_a, "pp0"_a, "pp1"_a, "options"_a, "est_config"_a = EstimatorConfig(),
          "devices"_a = std::vector<int>());
    m.def("HybridEstimatePoseScaleOffsetTwoFocalBatch", &estimate_many<PoseScaleOffsetTwoFocal, MP_KIND_TWO_FOCAL>, "x0"_a,
          "x1"_a, "depth0"_a, "depth1"_a, "min_depth"_a, "pp0"_a, "pp1"_a, "options"_a, "est_config"_a = EstimatorConfig(),
          "devices"_a = std::vector<int>());
    m.def("HybridEstimatePoseAndScale", &estimate_scale_only, "x0"_a, "x1"_a, "depth0"_a, "depth1"_a, "K0"_a, "K1"_a, "options"_a,
          "est_config"_a = EstimatorConfig());
    m.def("HybridEstimatePoseAndScaleBatch", &estimate_scale_only_many, "x0"_a, "x1"_a, "depth0"_a, "depth1"_a, "K0"_a, "K1"_a,
          "options"_a, "est_config"_a = EstimatorConfig(), "devices"_a = std::vector<int>());
    m.def("device_count", &mp_device_count, "CUDA devices visible to the process");
    m.def("estimate_scale_and_pose", &scale_and_pose, "X"_a, "Y"_a, "W"_a);
    m.def("solve_scale_and_shift", &solve_raw<MP_KIND_CALIBRATED>, "x_homo"_a, "y_homo"_a, "depth_x"_a, "depth_y"_a);
    m.def("solve_scale_and_shift_shared_focal", &solve_raw<MP_KIND_SHARED_FOCAL>, "x_homo"_a, "y_homo"_a, "depth_x"_a, "depth_y"_a);
    m.def("solve_scale_and_shift_two_focal", &solve_raw<MP_KIND_TWO_FOCAL>, "x_homo"_a, "y_homo"_a, "depth_x"_a, "depth_y"_a);
    m.def("solve_scale_shift_pose", &solve_pose<PoseScaleOffset, MP_KIND_CALIBRATED>, "x_homo"_a, "y_homo"_a, "depth_x"_a, "depth_y"_a);
    m.def("solve_scale_shift_pose_shared_focal", &solve_pose<PoseScaleOffsetSharedFocal, MP_KIND_SHARED_FOCAL>, "x_homo"_a, "y_homo"_a,
          "depth_x"_a, "depth_y"_a);
    m.def("solve_scale_shift_pose_two_focal", &solve_pose<PoseScaleOffsetTwoFocal, MP_KIND_TWO_FOCAL>, "x_homo"_a, "y_homo"_a,
          "depth_x"_a, "depth_y"_a);
}
