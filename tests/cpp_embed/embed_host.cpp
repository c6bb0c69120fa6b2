// A C++ host that drives the Python mirror the way the reference's C++ does: an embedded CPython, objects created on
// the main thread, the GIL released, and two worker threads ("tracking" and "local mapping") that re-take it around
// every call and move data across as column-major float matrices (what pybind11's Eigen casters hand over).
// Call sites mirrored (reference paths):  src/System.cc:91-99,158 (interpreter, get_configs / get_decoder /
// get_sequence, GIL release), include/System.h:57-71 (GIL guard), src/LocalMapping.cc:38-40 (Optimizer, MeshExtractor),
// src/Tracking_util.cc:35-49 (get_frame_by_id), src/LocalMapping_util.cc:94-111,165-206 (reconstruct_object,
// estimate_pose_cam_obj, extract_mesh_from_code).  Eigen is not in this image: ColMajor below stands in for
// Eigen::MatrixXf and the arrays are built with pybind11's f_style flag, which is the layout the Eigen caster produces.
//
//   embed_host --dry <pkg_dir>                        interpreter + GIL hand-over + array round trip, no GPU
//   embed_host <pkg_dir> <sequence_dir> <config.json> the full sequence on cuda:0
#include <pybind11/embed.h>
#include <pybind11/numpy.h>

#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace py = pybind11;
using FArray = py::array_t<float, py::array::f_style | py::array::forcecast>;

struct ColMajor {                       // rows x cols, column-major, owned by C++ (no Python object behind it)
    int rows = 0, cols = 0;
    std::vector<float> v;
};

struct GilGuard {                       // what the reference calls PyThreadStateLock
    PyGILState_STATE st;
    GilGuard() : st(PyGILState_Ensure()) {}
    ~GilGuard() { PyGILState_Release(st); }
};

static ColMajor to_cpp(const py::handle& obj) {          // ".cast<Eigen::MatrixXf>()"
    FArray a = FArray::ensure(obj);
    if (!a) throw std::runtime_error("not convertible to a float matrix");
    ColMajor m;
    m.rows = (int)a.shape(0);
    m.cols = a.ndim() > 1 ? (int)a.shape(1) : 1;
    m.v.assign(a.data(), a.data() + (size_t)m.rows * m.cols);
    return m;
}

static FArray to_py(const ColMajor& m, bool vector = false) {   // Eigen matrix -> numpy (a copy, Fortran order)
    FArray a = vector ? FArray({(py::ssize_t)m.rows}) : FArray({(py::ssize_t)m.rows, (py::ssize_t)m.cols});
    std::memcpy(a.mutable_data(), m.v.data(), m.v.size() * sizeof(float));
    return a;
}

struct Detection { ColMajor T_cam_obj, pts, rays, depth; bool has_rays = false; };

struct Shared {
    std::mutex mu;
    std::condition_variable cv;
    std::deque<Detection> queue;
    bool done = false;
    int n_det = 0, n_good = 0, n_faces = 0;
    std::string error;
};

int main(int argc, char** argv) {
    const bool dry = argc >= 3 && std::string(argv[1]) == "--dry";
    if (!dry && argc < 4) { std::fprintf(stderr, "usage: embed_host [--dry] <pkg_dir> [<sequence_dir> <config.json>]\n"); return 2; }
    const std::string pkg = dry ? argv[2] : argv[1];
    Shared sh;
    py::initialize_interpreter();
    py::object pyCfg, pyDecoder, pySequence, pyOptimizer, pyMeshExtractor;
    try {
        py::module_::import("sys").attr("path").attr("insert")(0, pkg);
        py::module_ io_utils = py::module_::import("reconstruct.utils");
        if (!dry) {
            pyCfg = io_utils.attr("get_configs")(std::string(argv[3]));
            pyDecoder = io_utils.attr("get_decoder")(pyCfg);
            pySequence = py::module_::import("reconstruct").attr("get_sequence")(std::string(argv[2]), pyCfg);
            py::module_ optim = py::module_::import("reconstruct.optimizer");
            pyOptimizer = optim.attr("Optimizer")(pyDecoder, pyCfg);
            pyMeshExtractor = optim.attr("MeshExtractor")(pyDecoder, pyCfg.attr("optimizer").attr("code_len"), pyCfg.attr("voxels_dim"));
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "set-up failed: %s\n", e.what());
        return 1;
    }
    PyThreadState* main_ts = PyEval_SaveThread();       // the main thread lets go of the GIL for good (System.cc:158)

    std::thread tracking([&] {
        try {
            GilGuard gil;
            if (dry) {                                   // a 5x3 matrix made in Python, taken over by C++
                py::object np = py::module_::import("numpy");
                Detection d;
                d.pts = to_cpp(np.attr("arange")(15).attr("reshape")(5, 3).attr("astype")("float32"));
                std::lock_guard<std::mutex> lk(sh.mu);
                sh.queue.push_back(std::move(d));
            } else {
                py::list dets = pySequence.attr("get_frame_by_id")(0);
                for (py::handle det : dets) {
                    Detection d;
                    d.T_cam_obj = to_cpp(det.attr("T_cam_obj"));
                    d.pts = to_cpp(det.attr("surface_points"));
                    if (!det.attr("rays").is_none()) {
                        d.rays = to_cpp(det.attr("rays"));
                        d.depth = to_cpp(det.attr("depth"));
                        d.has_rays = true;
                    }
                    std::lock_guard<std::mutex> lk(sh.mu);
                    sh.queue.push_back(std::move(d));
                    sh.cv.notify_one();
                }
            }
        } catch (const std::exception& e) { std::lock_guard<std::mutex> lk(sh.mu); sh.error = std::string("tracking: ") + e.what(); }
        std::lock_guard<std::mutex> lk(sh.mu);
        sh.done = true;
        sh.cv.notify_all();
    });

    std::thread mapping([&] {
        for (;;) {
            Detection d;
            {
                std::unique_lock<std::mutex> lk(sh.mu);  // waits WITHOUT the GIL, as LocalMapping::Run does
                sh.cv.wait(lk, [&] { return !sh.queue.empty() || sh.done; });
                if (sh.queue.empty()) return;
                d = std::move(sh.queue.front());
                sh.queue.pop_front();
            }
            try {
                GilGuard gil;
                ++sh.n_det;
                if (dry) {                               // back into Python: column-major copy, values intact
                    FArray a = to_py(d.pts);
                    const float s = a.attr("sum")().cast<float>(), a41 = a.attr("__getitem__")(py::make_tuple(4, 1)).cast<float>();
                    if (s != 105.f || a41 != 13.f || !a.attr("flags").attr("f_contiguous").cast<bool>()) throw std::runtime_error("round trip changed the matrix");
                    ++sh.n_good;
                    continue;
                }
                if (!d.has_rays || d.pts.rows < 50) continue;                       // LocalMapping_util.cc:170
                py::object mo = pyOptimizer.attr("reconstruct_object")(to_py(d.T_cam_obj), to_py(d.pts), to_py(d.rays), to_py(d.depth, true));
                if (!mo.attr("is_good").cast<bool>()) continue;
                ColMajor Two = to_cpp(mo.attr("t_cam_obj")), code = to_cpp(mo.attr("code"));
                if (Two.rows != 4 || Two.cols != 4 || code.rows != pyOptimizer.attr("code_len").cast<int>()) throw std::runtime_error("unexpected result shapes");
                py::object mesh = pyMeshExtractor.attr("extract_mesh_from_code")(to_py(code, true));
                ColMajor verts = to_cpp(mesh.attr("vertices"));
                py::array_t<int, py::array::f_style | py::array::forcecast> faces(mesh.attr("faces"));
                if (verts.cols != 3 || faces.shape(1) != 3 || faces.shape(0) == 0) throw std::runtime_error("empty mesh");
                // re-observation: pose-only update from the scale-free pose (LocalMapping_util.cc:94-111)
                double det3 = 0;
                { const float* m = Two.v.data(); auto at = [&](int r, int c) { return (double)m[c * 4 + r]; };
                  det3 = at(0,0)*(at(1,1)*at(2,2)-at(1,2)*at(2,1)) - at(0,1)*(at(1,0)*at(2,2)-at(1,2)*at(2,0)) + at(0,2)*(at(1,0)*at(2,1)-at(1,1)*at(2,0)); }
                const float s = (float)std::cbrt(det3);
                ColMajor SE3 = Two;
                for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) SE3.v[c * 4 + r] /= s;
                ColMajor T2 = to_cpp(pyOptimizer.attr("estimate_pose_cam_obj")(to_py(SE3), s, to_py(d.pts), to_py(code, true)));
                for (float x : T2.v) if (!(x == x)) throw std::runtime_error("NaN pose");
                ++sh.n_good;
                sh.n_faces += (int)faces.shape(0);
            } catch (const std::exception& e) { std::lock_guard<std::mutex> lk(sh.mu); sh.error = std::string("mapping: ") + e.what(); return; }
        }
    });

    tracking.join();
    mapping.join();
    PyEval_RestoreThread(main_ts);
    pyMeshExtractor = py::object(); pyOptimizer = py::object(); pySequence = py::object(); pyDecoder = py::object(); pyCfg = py::object();
    if (!sh.error.empty()) { std::fprintf(stderr, "%s\n", sh.error.c_str()); return 1; }
    std::printf("embed host ok: detections=%d good=%d faces=%d\n", sh.n_det, sh.n_good, sh.n_faces);
    std::fflush(stdout);
    if (dry) py::finalize_interpreter();   // full mode: leave teardown to process exit (torch's CUDA state outlives Py_Finalize badly)
    return 0;
}
