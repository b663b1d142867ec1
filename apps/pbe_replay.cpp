// pbe_replay -- Foam-free replay of the reference's "run only the PBE" driver
// (applications/PBEFoam/PBEFoam.C:56-67: while (runTime.run()) { runTime++; fluid.correct(); })
// on synthetic per-cell fields: BASELINE.json configs[4].  MOC and QMOM run side by side through
// the PBEMethod mirror of include/pbe_method.hpp; every step does what PBEDiameter::correct() does
// (PBEDiameter.H:95-99): method_->correct(); d_ = method_->d().
//
//   pbe_replay [--cells 4000000] [--steps 10] [--classes 50] [--order 3] [--device 0]
//              [--ranks N]          one process per GPU (forked here, devices device..device+N-1): the cells are
//                                   block-partitioned over the ranks (BASELINE configs[3]), an NCCL communicator
//                                   is built through the C ABI (pbe_comm_unique_id / pbe_comm_init) and every
//                                   printed diagnostic is the global one (pbe_reduce_diagnostics*)
//              [--dict file]        (a PBEDiameterCoeffs-style dictionary; default: the c01 case)
//              [--mesh nx ny nz [--velocity u v w]]   hex block in OpenFOAM face addressing: both methods then
//                                   run solveWithMULES (source + explicit upwind transport, MOC.C:198-272,
//                                   QMOM.C:131-197)
//              [--implicit [--courant C]]  with --mesh: usingMULES stays at the reference's default (false) and both
//                                   methods run solveWithFVM (fvm::ddt + fvm::div == S, MOC.C:186-198, QMOM.C:201-209),
//                                   fvSolution controls of the reference's cases (tolerance 1e-5, relTol 0, minIter 1);
//                                   --courant sets deltaT from the block's largest cell Courant number
//
// Prints one JSON line per method per step (wall ms with the diameter read-back, volume-averaged
// / max / min diameter as printAvgMaxMin would, Utility.C:4-10) and a summary line per method.
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <iterator>
#include <thread>

#include "pbe_method.hpp"

namespace {

// splitmix64: small, seedable, identical on every host
struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    double uniform() { return (double)(next() >> 11) * (1.0 / 9007199254740992.0); }
    double uniform(double a, double b) { return a + (b - a) * uniform(); }
};

// c01 validation case (tests/validation/simmons_and_azzopardi/c01/constant/phaseProperties:24-50,
// .../thermophysicalProperties.acps:38,47, .../system/controlDict:28)
const char *kDefaultDict = R"(
    coalescenceKernel CoulaloglouTavlaridesC;
    breakupKernel CoulaloglouTavlarides;
    daughterParticleDistribution uniformBinaryBreakup;
    CoulaloglouTavlaridesCCoeffs { c1 4.5e6; c2 1.891e12; gamma 0; sigma 0.01; }
    CoulaloglouTavlaridesCoeffs  { c1 0.1;   c2 0.06354;  gamma 0; sigma 0.01; }
    uniformBinaryBreakupCoeffs { }
)";

double lognormalVolumePdf(double v, double dmean = 5.0e-4, double dstd = 5.0e-5) {
    // number pdf in volume space of a log-normal diameter distribution
    // (tests/validation/simmons_and_azzopardi/setupCases.py:22-40)
    const double var = dstd * dstd;
    const double mu = std::log(dmean / std::sqrt(1 + var / (dmean * dmean)));
    const double sg = std::sqrt(std::log(1 + var / (dmean * dmean)));
    const double d = std::cbrt(6.0 * v / M_PI);
    const double z = (std::log(d) - mu) / sg;
    return std::exp(-0.5 * z * z) / (d * sg * std::sqrt(2 * M_PI)) * (2.0 / (M_PI * d * d));
}

struct Stats {
    double avg, mx, mn;
};
// printAvgMaxMin (Utility.C:4-10): with ranks the three numbers are reduced over them through the C ABI
Stats avgMaxMin(pbe_ctx *ctx, const std::vector<double> &f, const std::vector<double> &V) {
    double sums[2] = {0, 0}, mx = -INFINITY, mn = INFINITY;
    for (size_t c = 0; c < f.size(); ++c) {
        sums[0] += f[c] * V[c];
        sums[1] += V[c];
        mx = std::max(mx, f[c]);
        mn = std::min(mn, f[c]);
    }
    pbe::check(pbe_reduce_diagnostics_host(ctx, 2, sums, 1, &mn, 1, &mx), "printAvgMaxMin");
    return {sums[0] / sums[1], mx, mn};
}

// rank, ranks and the barrier-ish max over the ranks of a scalar (the slowest rank's time is the step's)
struct Ranks {
    int n = 1, rank = 0;
    pbe_ctx *ctx = nullptr;
    double maxOver(double x) const {
        pbe::check(pbe_reduce_diagnostics_host(ctx, 0, nullptr, 0, nullptr, 1, &x), "max over ranks");
        return x;
    }
};

void replay(const char *label, pbe::PBEMethod &m, const Ranks &R, const std::vector<double> &V, int steps,
            int64_t globalCells, double unitsPerCell, const char *unit, double dScale = 1.0) {
    double total = 0;
    for (int s = 0; s < steps; ++s) {
        (void)R.maxOver(0.0); // the ranks start the step together
        auto t0 = std::chrono::steady_clock::now();
        m.correct();                              // PBEDiameter::correct(): method_->correct();
        const std::vector<double> &d = m.d();     //                         d_ = method_->d();
        auto t1 = std::chrono::steady_clock::now();
        const double ms = R.maxOver(std::chrono::duration<double, std::milli>(t1 - t0).count());
        if (s > 0) total += ms; // step 0 carries one-time warm-up
        Stats st = avgMaxMin(R.ctx, d, V);
        st.avg *= dScale, st.mx *= dScale, st.mn *= dScale;
        if (R.rank == 0)
            std::printf("{\"method\": \"%s\", \"step\": %d, \"ms\": %.3f, \"d_avg\": %.9e, \"d_max\": %.9e, "
                        "\"d_min\": %.9e}\n",
                        label, s, ms, st.avg, st.mx, st.mn);
    }
    if (steps > 1 && R.rank == 0) {
        const double ms = total / (steps - 1);
        std::printf("{\"method\": \"%s\", \"summary\": true, \"ranks\": %d, \"cells\": %lld, \"ms_per_step\": %.3f, "
                    "\"%s\": %.4e}\n",
                    label, R.n, (long long)globalCells, ms, unit, globalCells * unitsPerCell / (ms * 1e-3));
    }
}

} // namespace

int main(int argc, char **argv) {
    int64_t cells = 4000000;
    int steps = 10, classes = 50, order = 3, device = 0, ranks = 1;
    std::string dictFile;
    int meshN[3] = {0, 0, 0};
    double vel[3] = {1.0, 0.5, 0.25};
    bool implicitSolve = false;
    double courant = 0.0;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto val = [&]() -> const char * {
            if (i + 1 >= argc) {
                std::fprintf(stderr, "missing value after %s\n", a.c_str());
                std::exit(2);
            }
            return argv[++i];
        };
        if (a == "--cells") cells = (int64_t)std::atof(val());
        else if (a == "--steps") steps = std::atoi(val());
        else if (a == "--classes") classes = std::atoi(val());
        else if (a == "--order") order = std::atoi(val());
        else if (a == "--device") device = std::atoi(val());
        else if (a == "--ranks") ranks = std::atoi(val());
        else if (a == "--dict") dictFile = val();
        else if (a == "--mesh") { // nx ny nz hex block: MOC runs solveWithMULES (transport + source)
            for (int k = 0; k < 3; ++k) meshN[k] = std::atoi(val());
            cells = (int64_t)meshN[0] * meshN[1] * meshN[2];
        } else if (a == "--implicit")
            implicitSolve = true;
        else if (a == "--courant")
            courant = std::atof(val());
        else if (a == "--velocity")
            for (int k = 0; k < 3; ++k) vel[k] = std::atof(val());
        else {
            std::fprintf(stderr, "unknown option %s\n", a.c_str());
            return 2;
        }
    }
    // ---- one process per GPU: fork BEFORE anything touches CUDA; the children are the ranks ---------
    Ranks R;
    std::string idFile;
    if (ranks > 1) {
        if (meshN[0] > 0) {
            std::fprintf(stderr, "--ranks with --mesh: the decomposed transport (processor patches + pbe_halo_exchange) is "
                                 "driven from openpbe_b200/sharding.py; the replay driver shards the source path only\n");
            return 2;
        }
        idFile = "/tmp/pbe_replay_nccl_id." + std::to_string((long)getpid());
        std::vector<pid_t> kids;
        int myRank = -1;
        for (int r = 0; r < ranks; ++r) {
            const pid_t pid = fork();
            if (pid == 0) {
                myRank = r;
                break;
            }
            kids.push_back(pid);
        }
        if (myRank < 0) { // the launcher: wait for the ranks
            int worst = 0;
            for (pid_t k : kids) {
                int st = 0;
                waitpid(k, &st, 0);
                if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) worst = 1;
            }
            std::remove(idFile.c_str());
            return worst;
        }
        R.n = ranks;
        R.rank = myRank;
        device += myRank;
    }
    const int64_t globalCells = cells;
    if (R.n > 1) { // contiguous, balanced block of the cell range (cells are independent: no data-path collective)
        const int64_t base = cells / R.n, rem = cells % R.n;
        cells = base + (R.rank < rem ? 1 : 0);
    }
    try {
        pbe::Context ctx(device);
        R.ctx = ctx.get();
        if (R.n > 1) {
            char id[PBE_COMM_UNIQUE_ID_BYTES];
            if (R.rank == 0) {
                pbe::check(pbe_comm_unique_id(id), "pbe_comm_unique_id");
                const std::string tmp = idFile + ".tmp";
                std::FILE *f = std::fopen(tmp.c_str(), "wb");
                if (!f || std::fwrite(id, 1, sizeof(id), f) != sizeof(id)) throw pbe::FatalError("cannot write " + tmp);
                std::fclose(f);
                std::rename(tmp.c_str(), idFile.c_str());
            } else {
                std::FILE *f = nullptr;
                for (int tries = 0; tries < 6000 && !(f = std::fopen(idFile.c_str(), "rb")); ++tries)
                    std::this_thread::sleep_for(std::chrono::milliseconds(10));
                if (!f || std::fread(id, 1, sizeof(id), f) != sizeof(id)) throw pbe::FatalError("no NCCL id from rank 0");
                std::fclose(f);
            }
            pbe::check(pbe_comm_init(ctx.get(), R.n, R.rank, id), "pbe_comm_init");
        }
        std::string text = kDefaultDict;
        if (!dictFile.empty()) {
            std::ifstream in(dictFile);
            if (!in) throw pbe::FatalError("cannot open " + dictFile);
            text.assign(std::istreambuf_iterator<char>(in), std::istreambuf_iterator<char>());
        }
        // ---- synthetic phase fields, seed 20240611 -------------------------------
        Rng rng(20240611 + 7919ull * (uint64_t)R.rank);
        std::vector<double> eps(cells), alpha(cells), rho(cells, 1166.0), nu(cells, 1.599752e-3 / 1166.0),
            V(cells, 1e-9);
        for (int64_t c = 0; c < cells; ++c) {
            eps[c] = std::pow(10.0, rng.uniform(-3.0, 0.0));
            alpha[c] = rng.uniform(0.01, 0.10);
        }
        pbe::Phase phase;
        phase.name = "acps";
        phase.ncells = cells;
        phase.alpha = alpha.data();
        phase.rho = rho.data();
        phase.nu = nu.data();
        phase.epsilon = eps.data();
        phase.V = V.data();
        phase.deltaT = 2.5e-4;
        phase.device = device;
        phase.ctx = ctx.get(); // both methods share the rank's context (and its communicator)
        // ---- optional hex block in OpenFOAM face addressing (blockMesh order: internal faces by
        // owner then neighbour; patches x-, x+, y-, y+, z-, z+), unit cube, uniform velocity ----
        pbe::FaceMesh fm;
        std::vector<int32_t> own, nei, bown;
        std::vector<double> phiF;
        if (meshN[0] > 0) {
            const int nx = meshN[0], ny = meshN[1], nz = meshN[2];
            const double dx = 1.0 / nx, dy = 1.0 / ny, dz = 1.0 / nz;
            std::fill(V.begin(), V.end(), dx * dy * dz);
            for (int64_t c = 0; c < cells; ++c) alpha[c] *= 1e-9 / (dx * dy * dz); // n0 ~ alpha V as without a mesh
            auto id = [&](int i, int j, int k) { return (int32_t)(i + (int64_t)nx * (j + (int64_t)ny * k)); };
            for (int k = 0; k < nz; ++k)
                for (int j = 0; j < ny; ++j)
                    for (int i = 0; i < nx; ++i) {
                        if (i + 1 < nx) own.push_back(id(i, j, k)), nei.push_back(id(i + 1, j, k)), phiF.push_back(vel[0] * dy * dz);
                        if (j + 1 < ny) own.push_back(id(i, j, k)), nei.push_back(id(i, j + 1, k)), phiF.push_back(vel[1] * dx * dz);
                        if (k + 1 < nz) own.push_back(id(i, j, k)), nei.push_back(id(i, j, k + 1)), phiF.push_back(vel[2] * dx * dy);
                    }
            auto patch = [&](int axis, int side) {
                const int n[3] = {nx, ny, nz};
                const double area[3] = {dy * dz, dx * dz, dx * dy};
                int lo[3] = {0, 0, 0}, hi[3] = {nx, ny, nz};
                lo[axis] = side ? n[axis] - 1 : 0;
                hi[axis] = lo[axis] + 1;
                for (int k = lo[2]; k < hi[2]; ++k)
                    for (int j = lo[1]; j < hi[1]; ++j)
                        for (int i = lo[0]; i < hi[0]; ++i) {
                            bown.push_back(id(i, j, k));
                            phiF.push_back((side ? 1.0 : -1.0) * vel[axis] * area[axis]);
                        }
            };
            for (int axis = 0; axis < 3; ++axis) patch(axis, 0), patch(axis, 1);
            fm.nInternalFaces = (int64_t)own.size();
            fm.nBoundaryFaces = (int64_t)bown.size();
            fm.owner = own.data(), fm.neighbour = nei.data(), fm.boundaryOwner = bown.data();
            fm.phi = phiF.data();
            phase.mesh = &fm;
            // CFL 0.4 on the block
            phase.deltaT = 0.4 / (std::fabs(vel[0]) * nx + std::fabs(vel[1]) * ny + std::fabs(vel[2]) * nz + 1e-300);
            if (phase.deltaT > 2.5e-4) phase.deltaT = 2.5e-4;
            if (courant > 0.0) // the implicit solve is not bound by the Courant number: let the caller pick it
                phase.deltaT = courant / (std::fabs(vel[0]) * nx + std::fabs(vel[1]) * ny + std::fabs(vel[2]) * nz + 1e-300);
        }

        // QMOM starts from the moments of the MOC initial condition (same physical state in both
        // methods, as the reference's side-by-side validation cases do:
        // tests/integration/QMOM/setupCases.py:44-48 integrates the same pdf MOC discretises)
        //
        // QMOM clamps every moment to OpenFOAM's SMALL = 1e-15 after the update (QMOM.C:117-123),
        // so with volumes in m^3 (xi ~ 1e-10) every moment above m1 would be overwritten.  The
        // reference only ever runs QMOM with O(1) volumes (tests/integration/QMOM/setupCases.py:
        // 44-48, v0 = 0.5), so the replay does the same: volumes in units of u = vmax of the class
        // grid, the Coulaloglou-Tavlarides constants carried over through the kernels'
        // homogeneity in xi (c1 u^(7/9), c2 u^(4/3) for coalescence; c1 u^(-2/9), c2 u^(-5/9) for
        // breakup: identical physical rates), diameters scaled back by u^(1/3) when printed.
        std::vector<double> m0((size_t)2 * order * cells, 0.0);
        double volUnit = 1.0;
        {   // ---- methodOfClasses: c01 grid, xi1 = 3.5918876006043295e-12 at 50 classes -----------
            pbe::Dictionary dict = pbe::Dictionary::parse(text, "PBEDiameterCoeffs");
            dict.set("method", "methodOfClasses");
            const double xi1 = 3.5918876006043295e-12 * 50.0 / classes;
            dict.subDictOrAdd("MOCCoeffs").set("numberOfClasses", (double)classes).set("xi1", xi1);
            if (phase.mesh && !implicitSolve) dict.subDictOrAdd("MOCCoeffs").set("usingMULES", std::string("true"));
            std::vector<double> p(classes), n0((size_t)classes * cells);
            double norm = 0;
            for (int i = 0; i < classes; ++i) {
                p[i] = lognormalVolumePdf(xi1 * (i + 1)) + 1e-300;
                norm += p[i] * xi1 * (i + 1);
            }
            for (int i = 0; i < classes; ++i)
                for (int64_t c = 0; c < cells; ++c)
                    n0[(size_t)i * cells + c] = alpha[c] * V[c] * p[i] / norm * std::exp2(rng.uniform(-1.0, 1.0));
            volUnit = xi1 * classes;
            for (int i = 0; i < classes; ++i) {
                const double xi = xi1 * (i + 1) / volUnit;
                double pw = 1.0;
                for (int k = 0; k < 2 * order; ++k, pw *= xi)
                    for (int64_t c = 0; c < cells; ++c) m0[(size_t)k * cells + c] += pw * n0[(size_t)i * cells + c];
            }
            pbe::PBEMethods::MOC moc(dict, phase, n0.data());
            std::vector<double>().swap(n0);
            replay("methodOfClasses", moc, R, V, steps, globalCells, classes, "cell_class_updates_per_s");
            if (R.rank == 0) // global values: the library reduced them over the ranks inside correct()
                std::printf("{\"method\": \"methodOfClasses\", \"min_n0\": %.6e, \"continuity_error\": %.6e}\n",
                            moc.classMin(0), moc.continuityError());
            if (R.rank == 0 && phase.mesh && implicitSolve) { // OpenFOAM's "Solving for n<i>, ..." lines, condensed
                int itMin = 1 << 30, itMax = 0;
                double r0 = 0, r1 = 0;
                for (const pbe_fvm_perf &q : moc.solverPerformance()) {
                    itMin = std::min(itMin, q.iterations), itMax = std::max(itMax, q.iterations);
                    r0 = std::max(r0, q.initial_residual), r1 = std::max(r1, q.final_residual);
                }
                std::printf("{\"method\": \"methodOfClasses\", \"transport\": \"solveWithFVM\", \"iterations\": [%d, %d], "
                            "\"initial_residual_max\": %.3e, \"final_residual_max\": %.3e}\n", itMin, itMax, r0, r1);
            }
        }
        {   // ---- quadratureMethodOfMoments: the 2*order moments of the same initial condition ------
            pbe::Dictionary dict = pbe::Dictionary::parse(text, "PBEDiameterCoeffs");
            dict.set("method", "quadratureMethodOfMoments");
            dict.subDictOrAdd("QMOMCoeffs").set("quadratureOrder", (double)order);
            auto rescale = [&](const char *sub, double p1, double p2, double d1, double d2) {
                if (!dict.found(sub)) return;
                pbe::Dictionary &cd = dict.subDictOrAdd(sub);
                cd.set("c1", cd.lookupOrDefault<double>("c1", d1) * std::pow(volUnit, p1));
                cd.set("c2", cd.lookupOrDefault<double>("c2", d2) * std::pow(volUnit, p2));
            };
            rescale("CoulaloglouTavlaridesCCoeffs", 7.0 / 9.0, 4.0 / 3.0, 0.00487, 0.008);
            rescale("CoulaloglouTavlaridesCoeffs", -2.0 / 9.0, -5.0 / 9.0, 0.00487, 0.008);
            if (phase.mesh && !implicitSolve) dict.subDictOrAdd("QMOMCoeffs").set("usingMULES", std::string("true"));
            pbe::PBEMethods::QMOM q(dict, phase, m0.data());
            std::vector<double>().swap(m0);
            replay("quadratureMethodOfMoments", q, R, V, steps, globalCells, 1.0, "cell_updates_per_s", std::cbrt(volUnit));
            const auto a = q.avgMaxMin(0);
            if (R.rank == 0)
                std::printf("{\"method\": \"quadratureMethodOfMoments\", \"m0_avg\": %.6e, \"m0_max\": %.6e, "
                            "\"m0_min\": %.6e}\n", a.avg, a.max, a.min);
        }
    } catch (const pbe::FatalError &e) {
        std::cerr << e.what() << "\n\nFOAM exiting\n";
        return 1;
    }
    return 0;
}
