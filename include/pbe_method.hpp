// pbe_method.hpp -- Foam-free C++ host mirror of OpenPBE's PBEMethod plug-in surface, on top of
// the C ABI of libpbe_b200.so (include/pbe_b200.h).  Header-only, C++17, no CUDA headers.
//
// What it mirrors (file:line relative to the reference root):
//   * class PBEMethod { virtual void correct() = 0; virtual const volScalarField d() const = 0;
//     static autoPtr<PBEMethod> New(const dictionary&, const phaseModel&); }   PBEMethod.H:62-126
//   * run-time selection by the dictionary word `method` with the reference's TypeNames
//     "methodOfClasses" (MOC.H:100), "quadratureMethodOfMoments" (QMOM.H:87), "methodOfMoments" (MOM.H:92), "dummy"
//     (dummy.H:67); unknown names fail listing the valid ones                  PBEMethod.C:81-108
//   * kernel selection by `coalescenceKernel` / `breakupKernel` /
//     `daughterParticleDistribution` with coefficients in `<TypeName>Coeffs`
//     (coalescenceKernel.C:46-71, breakupKernel.C:46-68, daughterParticleDistribution.C:54-81);
//     Coulaloglou-Tavlarides defaults c1 0.00487, c2 0.008, gamma 0, sigma 0.047
//     (coalescenceKernels/CoulaloglouTavlarides.C:39-42); constantCoalescence key `constant`
//     (constantCoalescence.C:48)
//   * MOCCoeffs { numberOfClasses; xi1; sourceRelaxation 1.0; usingMULES } (MOC.C:49-83),
//     QMOMCoeffs { quadratureOrder; usingMULES } (QMOM.C:59-82), dummyCoeffs { constantDiameter }
//   * errors: the reference aborts through FatalErrorIn(...) << exit(FatalError); here
//     pbe::FatalError is thrown with the same wording, and every non-zero C-ABI status is turned
//     into one.
//
// What it is not: OpenFOAM.  `Phase` stands in for phaseModel + fvMesh: the caller hands in the
// per-cell fields the kernels read (alpha, rho, nu, epsilon.<phase>, mesh.V()) and deltaT.  The
// mesh-coupled transport of MOC::solveWithMULES / QMOM::momentTransport is outside this path:
// without a FaceMesh correct() applies the zero-flux form  n += dt S  (MOC.C:259-272 with phi = 0), the
// form the reference's own 1-cell integration cases reduce to; with one it runs solveWithMULES on the
// device (Gauss upwind, or Gauss vanLeer with the MULES limiter).  The methods OWN their class / moment
// fields on the device (pbe_moc_state / pbe_qmom_state, as MOC.H:66); correct() sends this step's
// phase fields and brings d back (pbe_*_correct_resident); the field accessors copy back on demand.
// With Phase::ctx carrying a communicator (pbe_comm_init) the diagnostics are global.
#pragma once

#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "pbe_b200.h"

namespace pbe {

struct FatalError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void check(int rc, const char *where) {
    if (rc != PBE_OK) {
        std::ostringstream os;
        os << "--> PBE FATAL ERROR in " << where << " : " << pbe_last_error() << " (status " << rc << ")";
        throw FatalError(os.str());
    }
}

// ---- a minimal OpenFOAM-style dictionary: `key value;` entries and `name { ... }` sub-dicts -------
class Dictionary {
public:
    Dictionary() = default;
    explicit Dictionary(std::string name) : name_(std::move(name)) {}

    // parse `key value; sub { key value; }` text (comments // and /* */ are skipped)
    static Dictionary parse(const std::string &text, const std::string &name = "dictionary") {
        size_t pos = 0;
        Dictionary d(name);
        d.parseBody(strip(text), pos, false);
        return d;
    }

    Dictionary &set(const std::string &key, const std::string &value) {
        entries_[key] = value;
        return *this;
    }
    Dictionary &set(const std::string &key, double value) {
        std::ostringstream os;
        os.precision(17);
        os << value;
        return set(key, os.str());
    }
    Dictionary &subDictOrAdd(const std::string &key) {
        auto it = subs_.find(key);
        if (it == subs_.end()) it = subs_.emplace(key, Dictionary(name_ + "::" + key)).first;
        return it->second;
    }

    bool found(const std::string &key) const { return entries_.count(key) || subs_.count(key); }

    // dictionary::lookup: a missing keyword is fatal (OpenFOAM FatalIOError wording)
    const std::string &lookup(const std::string &key) const {
        auto it = entries_.find(key);
        if (it == entries_.end())
            throw FatalError("keyword " + key + " is undefined in dictionary \"" + name_ + "\"");
        return it->second;
    }
    template <class T>
    T lookupOrDefault(const std::string &key, const T &dflt) const {
        auto it = entries_.find(key);
        return it == entries_.end() ? dflt : convert<T>(it->second, key);
    }
    double readScalar(const std::string &key) const { return convert<double>(lookup(key), key); }
    int readLabel(const std::string &key) const { return convert<int>(lookup(key), key); }
    const Dictionary &subDict(const std::string &key) const {
        auto it = subs_.find(key);
        if (it == subs_.end())
            throw FatalError("keyword " + key + " is undefined in dictionary \"" + name_ + "\"");
        return it->second;
    }
    // <Type>Coeffs may be absent for coefficient-free kernels in hand-built dictionaries
    const Dictionary &optionalSubDict(const std::string &key) const {
        static const Dictionary empty("empty");
        auto it = subs_.find(key);
        return it == subs_.end() ? empty : it->second;
    }
    const std::string &name() const { return name_; }

private:
    template <class T>
    static T convert(const std::string &s, const std::string &key) {
        std::istringstream is(s);
        if constexpr (std::is_same<T, bool>::value) { // OpenFOAM Switch
            if (s == "on" || s == "true" || s == "yes" || s == "1") return true;
            if (s == "off" || s == "false" || s == "no" || s == "0") return false;
            throw FatalError("bad Switch value '" + s + "' for keyword " + key);
        } else {
            T v{};
            if (!(is >> v)) throw FatalError("cannot read value '" + s + "' of keyword " + key);
            return v;
        }
    }
    static std::string strip(const std::string &t) {
        std::string o;
        for (size_t i = 0; i < t.size(); ++i) {
            if (t.compare(i, 2, "//") == 0) {
                while (i < t.size() && t[i] != '\n') ++i;
            } else if (t.compare(i, 2, "/*") == 0) {
                size_t e = t.find("*/", i + 2);
                i = (e == std::string::npos) ? t.size() : e + 1;
            } else
                o += t[i];
        }
        return o;
    }
    static void skipWs(const std::string &t, size_t &p) {
        while (p < t.size() && std::isspace((unsigned char)t[p])) ++p;
    }
    static std::string word(const std::string &t, size_t &p) {
        skipWs(t, p);
        size_t b = p;
        while (p < t.size() && !std::isspace((unsigned char)t[p]) && t[p] != ';' && t[p] != '{' && t[p] != '}')
            ++p;
        return t.substr(b, p - b);
    }
    void parseBody(const std::string &t, size_t &p, bool nested) {
        for (;;) {
            skipWs(t, p);
            if (p >= t.size()) {
                if (nested) throw FatalError("unterminated '{' in dictionary " + name_);
                return;
            }
            if (t[p] == '}') {
                if (!nested) throw FatalError("unexpected '}' in dictionary " + name_);
                ++p;
                return;
            }
            const std::string key = word(t, p);
            if (key.empty()) throw FatalError("parse error in dictionary " + name_);
            skipWs(t, p);
            if (p < t.size() && t[p] == '{') {
                ++p;
                subDictOrAdd(key).parseBody(t, p, true);
            } else {
                size_t e = t.find(';', p);
                if (e == std::string::npos) throw FatalError("missing ';' after keyword " + key);
                std::string v = t.substr(p, e - p);
                while (!v.empty() && std::isspace((unsigned char)v.back())) v.pop_back();
                entries_[key] = v;
                p = e + 1;
            }
        }
    }
    std::string name_ = "dictionary";
    std::map<std::string, std::string> entries_;
    std::map<std::string, Dictionary> subs_;
};

// ---- stand-in for phaseModel + fvMesh: the per-cell fields the kernels read ----------------------
// All pointers are HOST arrays of ncells doubles owned by the caller and must outlive the method.
struct Phase {
    std::string name = "dispersed";
    int64_t ncells = 0;
    const double *alpha = nullptr;   // phase fraction (continuity check, MOC.C:254-279)
    const double *rho = nullptr;     // dispersed-phase density  (rhod_, CoulaloglouTavlarides.C:43)
    const double *nu = nullptr;      // dispersed-phase kinematic viscosity (nud_)
    const double *epsilon = nullptr; // "epsilon.<phase>" looked up from the mesh registry (:52-53)
    const double *V = nullptr;       // mesh.V()
    double deltaT = 0.0;             // runTime.deltaT()
    int device = 0;                  // CUDA device the state lives on
    const struct FaceMesh *mesh = nullptr; // fvMesh addressing + phase_.phi(): enables transport
    pbe_ctx *ctx = nullptr;          // optional: a context the caller owns (e.g. with a communicator); else one per method
};

// ---- stand-in for fvMesh's face addressing, phase_.phi() and fvSchemes' div(phi,n.*) entry ------------
// (what MOC::solveWithMULES reads, MOC.C:198-272).  HOST arrays owned by the caller; phi is re-read
// (uploaded) at every correct(), like the reference re-reads phase_.phi().
struct FaceMesh {
    int64_t nInternalFaces = 0, nBoundaryFaces = 0;
    const int32_t *owner = nullptr, *neighbour = nullptr; // fvMesh::owner() / neighbour()
    const int32_t *boundaryOwner = nullptr;               // faceCells of every patch, patch order
    const double *phi = nullptr;                          // [nInternalFaces + nBoundaryFaces]
    const double *boundaryValues = nullptr; // [fields][nBoundaryFaces] patch values, null = zeroGradient
    std::string divScheme = "Gauss upwind"; // fvSchemes: "div\\(phi,n.*\\)"; "Gauss vanLeer" needs the geometry:
    const double *Sf = nullptr;             // fvMesh::Sf() [nInternalFaces + nBoundaryFaces][3]
    const double *C = nullptr;              // fvMesh::C()  [ncells][3]
    const double *weights = nullptr;        // fvMesh::weights() [nInternalFaces]
    // usingMULES false -- the reference's default -- solves fvm::ddt + fvm::div(phi, n_k, "div(U,nk)") == S
    // (MOC.C:186-198, QMOM.C:201-209): the implicit term's scheme and what fvSolution says about the equation
    std::string implicitDivScheme = "Gauss upwind"; // fvSchemes: "div\\(U,nk\\)" / "div\\(phi.*,m.*\\)"
    double tolerance = 1e-5, relTol = 0.0;          // fvSolution: solvers { "n.*" { tolerance; relTol; minIter; maxIter; } }
    int minIter = 1, maxIter = 1000;                // (values of the reference's cases, e.g. tests/validation/.../c01/system/fvSolution:47-55)
    double relaxationFactor = 0.0;                  // fvSolution: relaxationFactors { equations { ".*" 1; } }; 0 = none
};

// ---- RAII over the C-ABI context and device buffers -------------------------------------------------
class Context {
public:
    explicit Context(int device) : owned_(true) { check(pbe_ctx_create(device, &h_), "pbe::Context"); }
    Context(int device, pbe_ctx *borrowed) : h_(borrowed), owned_(borrowed == nullptr) {
        if (!borrowed) check(pbe_ctx_create(device, &h_), "pbe::Context");
    }
    ~Context() {
        if (h_ && owned_) pbe_ctx_destroy(h_);
    }
    Context(const Context &) = delete;
    Context &operator=(const Context &) = delete;
    pbe_ctx *get() const { return h_; }

private:
    pbe_ctx *h_ = nullptr;
    bool owned_ = true;
};

class DeviceArray {
public:
    DeviceArray() = default;
    DeviceArray(Context &ctx, int64_t count) : ctx_(&ctx), count_(count) {
        if (count > 0) check(pbe_device_alloc(ctx.get(), count * (int64_t)sizeof(double), &p_), "pbe::DeviceArray");
    }
    ~DeviceArray() { reset(); }
    DeviceArray(DeviceArray &&o) noexcept { *this = std::move(o); }
    DeviceArray &operator=(DeviceArray &&o) noexcept {
        if (this != &o) {
            reset();
            ctx_ = o.ctx_, p_ = o.p_, count_ = o.count_;
            o.p_ = nullptr, o.count_ = 0;
        }
        return *this;
    }
    void upload(const double *host, int64_t count, int64_t offset = 0) {
        check(pbe_memcpy_h2d(ctx_->get(), data() + offset, host, count * (int64_t)sizeof(double)), "upload");
    }
    void download(double *host, int64_t count, int64_t offset = 0) const {
        check(pbe_memcpy_d2h(ctx_->get(), host, data() + offset, count * (int64_t)sizeof(double)), "download");
    }
    void zero() {
        if (count_) check(pbe_memset(ctx_->get(), p_, 0, count_ * (int64_t)sizeof(double)), "zero");
    }
    double *data() const { return static_cast<double *>(p_); }
    int64_t size() const { return count_; }

private:
    void reset() {
        if (p_) pbe_device_free(ctx_->get(), p_);
        p_ = nullptr;
    }
    Context *ctx_ = nullptr;
    void *p_ = nullptr;
    int64_t count_ = 0;
};

// ---- kernel selection: coalescenceKernel::New / breakupKernel::New /
// daughterParticleDistribution::New restated over the library's TypeName tables -----------------------
inline pbe_kernel_spec kernelSpecFromDict(const Dictionary &pbeDict) {
    pbe_kernel_spec s;
    std::memset(&s, 0, sizeof(s));
    const std::string ck = pbeDict.lookup("coalescenceKernel");
    check(pbe_coalescence_type(ck.c_str(), &s.coalescence), "coalescenceKernel::New");
    const Dictionary &cd = pbeDict.optionalSubDict(ck + "Coeffs");
    if (s.coalescence == PBE_COAL_CONSTANT) {
        s.coal_c[0] = cd.readScalar("constant"); // mandatory, constantCoalescence.C:48
    } else if (s.coalescence == PBE_COAL_CT) {
        s.coal_c[0] = cd.lookupOrDefault<double>("c1", 0.00487);
        s.coal_c[1] = cd.lookupOrDefault<double>("c2", 0.008);
        s.coal_c[2] = cd.lookupOrDefault<double>("gamma", 0.0);
        s.coal_c[3] = cd.lookupOrDefault<double>("sigma", 0.047);
    }
    const std::string bk = pbeDict.lookup("breakupKernel");
    check(pbe_breakup_type(bk.c_str(), &s.breakup), "breakupKernel::New");
    const Dictionary &bd = pbeDict.optionalSubDict(bk + "Coeffs");
    if (s.breakup == PBE_BR_CT) { // breakupKernels/CoulaloglouTavlarides.C:46-49
        s.br_c[0] = bd.lookupOrDefault<double>("c1", 0.00487);
        s.br_c[1] = bd.lookupOrDefault<double>("c2", 0.008);
        s.br_c[2] = bd.lookupOrDefault<double>("gamma", 0.0);
        s.br_c[3] = bd.lookupOrDefault<double>("sigma", 0.047);
    }
    const std::string dk = pbeDict.lookup("daughterParticleDistribution");
    check(pbe_daughter_type(dk.c_str(), &s.daughter), "daughterParticleDistribution::New");
    return s;
}

// ---- class PBEMethod (PBEMethod.H:62-126) ---------------------------------------------------------------
class PBEMethod {
public:
    using Factory = std::function<std::unique_ptr<PBEMethod>(const Dictionary &, const Phase &)>;

    virtual ~PBEMethod() = default;
    // one time step of the source path (the reference's correct() minus mesh transport)
    virtual void correct() = 0;
    // diameter field, one value per cell (host copy, refreshed lazily after correct())
    virtual const std::vector<double> &d() const = 0;

    // declareRunTimeSelectionTable / addToRunTimeSelectionTable
    static std::map<std::string, Factory> &table() {
        static std::map<std::string, Factory> t;
        return t;
    }
    static std::unique_ptr<PBEMethod> New(const Dictionary &pbeProperties, const Phase &phase);

protected:
    PBEMethod(const Dictionary &pbeProperties, const Phase &phase)
        : phase_(phase), ctx_(phase.device, phase.ctx), spec_(kernelSpecFromDict(pbeProperties)) {
        if (phase.ncells < 0 || phase.deltaT <= 0.0) throw FatalError("PBEMethod: bad phase (ncells, deltaT)");
    }
    // upload the per-cell kernel inputs once; update() refreshes them (coalescence_().update(), MOC.C:173)
    void stageCellFields() {
        const int64_t n = phase_.ncells;
        eps_ = DeviceArray(ctx_, n), rho_ = DeviceArray(ctx_, n), nu_ = DeviceArray(ctx_, n);
        if (phase_.alpha) alpha_ = DeviceArray(ctx_, n);
        if (phase_.V) vcell_ = DeviceArray(ctx_, n);
        updateCellFields();
    }

public:
    void updateCellFields() {
        const int64_t n = phase_.ncells;
        if (n == 0) return;
        if (!phase_.epsilon || !phase_.rho || !phase_.nu) throw FatalError("PBEMethod: phase lacks epsilon/rho/nu");
        eps_.upload(phase_.epsilon, n), rho_.upload(phase_.rho, n), nu_.upload(phase_.nu, n);
        if (phase_.alpha) alpha_.upload(phase_.alpha, n);
        if (phase_.V) vcell_.upload(phase_.V, n);
    }

protected:
    const Phase &phase_;
    Context ctx_;
    pbe_kernel_spec spec_;
    DeviceArray eps_, rho_, nu_, alpha_, vcell_;
};

namespace PBEMethods {

// page-lock a caller-owned host array so that correct() copies it asynchronously at full PCIe rate; a
// failure (already registered, not page-aligned enough for the driver) only costs speed
inline void tryPin(pbe_ctx *ctx, const void *p, int64_t bytes) {
    if (p && bytes > 0) (void)pbe_host_register(ctx, const_cast<void *>(p), bytes);
}
inline void tryUnpin(pbe_ctx *ctx, const void *p) {
    if (p) (void)pbe_host_unregister(ctx, const_cast<void *>(p));
}
// device mesh of a FaceMesh (+ geometry for the limited schemes)
inline pbe_fvm_controls fvmControls(const FaceMesh &fm) {
    pbe_fvm_controls c;
    c.tolerance = fm.tolerance, c.rel_tol = fm.relTol, c.min_iter = fm.minIter, c.max_iter = fm.maxIter;
    c.relaxation = fm.relaxationFactor, c.solver = PBE_FVM_SOLVER_AUTO;
    return c;
}

inline pbe_mesh *makeDeviceMesh(pbe_ctx *ctx, const Phase &phase, int scheme, const char *where) {
    const FaceMesh &fm = *phase.mesh;
    pbe_mesh *m = nullptr;
    check(pbe_mesh_create(ctx, phase.ncells, fm.nInternalFaces, fm.owner, fm.neighbour, fm.nBoundaryFaces, fm.boundaryOwner,
                          phase.V, &m),
          where);
    if (scheme != PBE_DIV_UPWIND) {
        if (!fm.Sf || !fm.C || !fm.weights) {
            pbe_mesh_destroy(m);
            throw FatalError(std::string(where) + ": the limited divergence scheme needs FaceMesh::Sf, C and weights");
        }
        check(pbe_mesh_set_geometry(m, fm.Sf, fm.C, fm.weights), where);
    }
    return m;
}

// ---- methodOfClasses (MOC.H:57-127, MOC.C:49-83, 159-184, 282-290) --------------------------------------
class MOC : public PBEMethod {
public:
    static constexpr const char *typeName = "methodOfClasses";

    // n0: initial class number fields n0..n{N-1}, HOST [N][ncells] class-major -- what the
    // reference reads as MUST_READ volScalarFields "n<i>" (MOC.C:69-77); may be null (zeros).
    MOC(const Dictionary &pbeProperties, const Phase &phase, const double *n0 = nullptr)
        : PBEMethod(pbeProperties, phase) {
        const Dictionary &c = pbeProperties.subDict("MOCCoeffs");
        N_ = c.readLabel("numberOfClasses");
        xi1_ = c.readScalar("xi1");
        usingMULES_ = c.lookupOrDefault<bool>("usingMULES", false);
        relax_ = c.lookupOrDefault<double>("sourceRelaxation", 1.0);
        // EXTENSION (no counterpart in the reference's MOCCoeffs): `classGridRatio r` > 1 makes the
        // class volumes geometric, xi_i = xi1 r^i, handled by the fixed-pivot technique (pbe_moc_create_pivots);
        // absent, the grid is the reference's uniform one, xi_i = (i + 1) xi1 (MOC.C:56,79)
        gridRatio_ = c.lookupOrDefault<double>("classGridRatio", 0.0);
        if (gridRatio_ > 0.0) {
            if (!(gridRatio_ > 1.0)) throw FatalError("MOC::MOC: classGridRatio must be > 1 (got " + std::to_string(gridRatio_) + ")");
            std::vector<double> piv((size_t)N_);
            for (int i = 0; i < N_; ++i) piv[(size_t)i] = (i == 0) ? xi1_ : piv[(size_t)i - 1] * gridRatio_;
            check(pbe_moc_create_pivots(ctx_.get(), N_, piv.data(), &spec_, relax_, &h_), "MOC::MOC");
        } else {
            check(pbe_moc_create(ctx_.get(), N_, xi1_, &spec_, relax_, &h_), "MOC::MOC");
        }
        const int64_t nc = phase.ncells;
        check(pbe_moc_state_create(h_, nc, phase.V, &state_), "MOC::MOC");
        if (n0) check(pbe_moc_state_upload(state_, nc, n0), "MOC::MOC");
        d_host_.assign((size_t)nc, 0.0);
        classMin_.assign((size_t)N_, 0.0);
        classMax_.assign((size_t)N_, 0.0);
        if (phase.mesh && nc) {
            if (!phase.V) throw FatalError("MOC: transport needs the cell volumes (Phase::V)");
            const std::string &sch = usingMULES_ ? phase.mesh->divScheme : phase.mesh->implicitDivScheme;
            check(pbe_div_scheme_type(sch.c_str(), &scheme_), usingMULES_ ? "MOC::MOC div(phi,n.*)" : "MOC::MOC div(U,nk)");
            mesh_ = makeDeviceMesh(ctx_.get(), phase, scheme_, "MOC::MOC mesh");
            check(pbe_moc_state_attach_mesh(state_, mesh_, scheme_), "MOC::MOC");
            if (!usingMULES_) { // solveWithFVM (MOC.C:177-180, 186-198), the reference's default
                const pbe_fvm_controls ctl = fvmControls(*phase.mesh);
                check(pbe_moc_state_set_implicit(state_, &ctl), "MOC::MOC usingMULES false");
            }
            tryPin(ctx_.get(), phase.mesh->phi, (phase.mesh->nInternalFaces + phase.mesh->nBoundaryFaces) * 8);
        }
        for (const double *p : {phase.epsilon, phase.rho, phase.nu, phase.alpha}) tryPin(ctx_.get(), p, nc * 8);
        tryPin(ctx_.get(), d_host_.data(), nc * 8);
        if (nc) check(pbe_moc_state_d(state_, d_host_.data()), "MOC::MOC d"); // PBEDiameter's d_(method_->d())
    }
    ~MOC() override {
        for (const double *p : {phase_.epsilon, phase_.rho, phase_.nu, phase_.alpha}) tryUnpin(ctx_.get(), p);
        if (phase_.mesh) tryUnpin(ctx_.get(), phase_.mesh->phi);
        tryUnpin(ctx_.get(), d_host_.data());
        if (state_) pbe_moc_state_destroy(state_);
        if (mesh_) pbe_mesh_destroy(mesh_);
        if (h_) pbe_moc_destroy(h_);
    }

    // MOC::correct() (MOC.C:159-184) + d() (PBEDiameter.H:95-99) in one library call: sources of every
    // class, limiter, solveWithMULES (or the zero-flux update), per-class min (the `Info << min(n_k)` of
    // MOC.C:181-183), continuity error (MOC.C:276-279), d -- the phase fields are re-read every step.
    void correct() override {
        if (phase_.ncells == 0) return;
        if (!phase_.epsilon || !phase_.rho || !phase_.nu) throw FatalError("PBEMethod: phase lacks epsilon/rho/nu");
        pbe_moc_diag diag;
        std::memset(&diag, 0, sizeof(diag));
        diag.class_min = classMin_.data();
        diag.class_max = classMax_.data();
        const FaceMesh *fm = mesh_ ? phase_.mesh : nullptr;
        check(pbe_moc_correct_resident(state_, phase_.epsilon, phase_.rho, phase_.nu, phase_.V ? phase_.alpha : nullptr,
                                       fm ? fm->phi : nullptr, fm ? fm->boundaryValues : nullptr,
                                       fm ? fm->nBoundaryFaces : 0, phase_.deltaT, d_host_.data(), &diag),
              "MOC::correct");
        continuityError_ = diag.cont_err_max;
    }
    const std::vector<double> &d() const override { return d_host_; }
    int numberOfClasses() const { return N_; }
    double xi(int i) const { // MOC.C:79, or the geometric extension
        return gridRatio_ > 0.0 ? xi1_ * std::pow(gridRatio_, i) : xi1_ * (i + 1);
    }
    // class field n<i>, host copy (runTime.write() of one field)
    std::vector<double> classNumberDensity(int i) const {
        const int64_t nc = phase_.ncells;
        std::vector<double> all((size_t)N_ * nc);
        if (nc) check(pbe_moc_state_download(state_, nc, all.data()), "MOC::classNumberDensity");
        return std::vector<double>(all.begin() + (size_t)i * nc, all.begin() + (size_t)(i + 1) * nc);
    }
    double classMin(int i) const { return classMin_[(size_t)i]; }
    double classMax(int i) const { return classMax_[(size_t)i]; }
    double continuityError() const { return continuityError_; }
    // "Solving for n<i>, Initial residual = .., Final residual = .., No Iterations .." of the last implicit step
    std::vector<pbe_fvm_perf> solverPerformance() const {
        std::vector<pbe_fvm_perf> p((size_t)N_);
        check(pbe_moc_state_solver_performance(state_, p.data()), "MOC::solverPerformance");
        return p;
    }
    double *deviceClassFields() const { // [N][ncells], for callers that chain their own device work
        double *n = nullptr;
        check(pbe_moc_state_device(state_, &n, nullptr, nullptr), "MOC::deviceClassFields");
        return n;
    }

private:
    pbe_moc *h_ = nullptr;
    pbe_moc_state *state_ = nullptr;
    int N_ = 0;
    double xi1_ = 0, relax_ = 1.0, continuityError_ = 0.0, gridRatio_ = 0.0;
    bool usingMULES_ = false;
    pbe_mesh *mesh_ = nullptr;
    int scheme_ = 0;
    std::vector<double> d_host_, classMin_, classMax_;
};

// ---- quadratureMethodOfMoments (QMOM.H:55-112, QMOM.C:59-129) ----------------------------------------------
class QMOM : public PBEMethod {
public:
    static constexpr const char *typeName = "quadratureMethodOfMoments";

    // m0: initial moment fields m0..m{2*order-1}, HOST [2*order][ncells] (MUST_READ "m<k>", QMOM.C:72-79)
    QMOM(const Dictionary &pbeProperties, const Phase &phase, const double *m0 = nullptr)
        : PBEMethod(pbeProperties, phase) {
        const Dictionary &c = pbeProperties.subDict("QMOMCoeffs");
        order_ = c.readLabel("quadratureOrder");
        usingMULES_ = c.lookupOrDefault<bool>("usingMULES", false);
        check(pbe_qmom_create(ctx_.get(), order_, &spec_, &h_), "QMOM::QMOM");
        const int64_t nc = phase.ncells;
        check(pbe_qmom_state_create(h_, nc, phase.V, &state_), "QMOM::QMOM");
        if (m0) check(pbe_qmom_state_upload(state_, nc, m0), "QMOM::QMOM");
        stats_.assign((size_t)(4 * order_ + 1) * 4, 0.0);
        if (phase.mesh && nc) {
            if (!phase.V) throw FatalError("QMOM: transport needs the cell volumes (Phase::V)");
            const std::string &sch = usingMULES_ ? phase.mesh->divScheme : phase.mesh->implicitDivScheme;
            check(pbe_div_scheme_type(sch.c_str(), &scheme_), "QMOM::QMOM div(phi,m.*)");
            mesh_ = makeDeviceMesh(ctx_.get(), phase, scheme_, "QMOM::QMOM mesh");
            check(pbe_qmom_state_attach_mesh(state_, mesh_, scheme_), "QMOM::QMOM");
            if (!usingMULES_) { // solveWithFVM (QMOM.C:112-115, 201-209), the reference's default
                const pbe_fvm_controls ctl = fvmControls(*phase.mesh);
                check(pbe_qmom_state_set_implicit(state_, &ctl), "QMOM::QMOM usingMULES false");
            }
            tryPin(ctx_.get(), phase.mesh->phi, (phase.mesh->nInternalFaces + phase.mesh->nBoundaryFaces) * 8);
        }
        d_host_.assign((size_t)nc, 0.0);
        for (const double *p : {phase.epsilon, phase.rho, phase.nu}) tryPin(ctx_.get(), p, nc * 8);
        tryPin(ctx_.get(), d_host_.data(), nc * 8);
        if (m0 && nc && order_ >= 1) check(pbe_qmom_state_d(state_, d_host_.data()), "QMOM::QMOM d"); // QMOM.C:81
    }
    ~QMOM() override {
        for (const double *p : {phase_.epsilon, phase_.rho, phase_.nu}) tryUnpin(ctx_.get(), p);
        if (phase_.mesh) tryUnpin(ctx_.get(), phase_.mesh->phi);
        tryUnpin(ctx_.get(), d_host_.data());
        if (state_) pbe_qmom_state_destroy(state_);
        if (mesh_) pbe_mesh_destroy(mesh_);
        if (h_) pbe_qmom_destroy(h_);
    }

    // QMOM::correct() (QMOM.C:85-129) in one library call: moment sources, NaN scrub, solveWithMULES (or the
    // zero-flux update), clamp to SMALL, d and the printAvgMaxMin partials of sources, moments and d
    void correct() override {
        if (phase_.ncells == 0) return;
        if (!phase_.epsilon || !phase_.rho || !phase_.nu) throw FatalError("PBEMethod: phase lacks epsilon/rho/nu");
        const int nm = 2 * order_;
        pbe_qmom_diag diag;
        diag.source = stats_.data();
        diag.moment = stats_.data() + (size_t)4 * nm;
        diag.d = stats_.data() + (size_t)8 * nm;
        const FaceMesh *fm = mesh_ ? phase_.mesh : nullptr;
        check(pbe_qmom_correct_resident(state_, phase_.epsilon, phase_.rho, phase_.nu, fm ? fm->phi : nullptr,
                                        fm ? fm->boundaryValues : nullptr, fm ? fm->nBoundaryFaces : 0, phase_.deltaT,
                                        d_host_.data(), &diag),
              "QMOM::correct");
    }
    const std::vector<double> &d() const override { return d_host_; }
    int quadratureOrder() const { return order_; }
    std::vector<double> moment(int k) const {
        const int64_t nc = phase_.ncells;
        std::vector<double> all((size_t)2 * order_ * nc);
        if (nc) check(pbe_qmom_state_download(state_, nc, all.data()), "QMOM::moment");
        return std::vector<double>(all.begin() + (size_t)k * nc, all.begin() + (size_t)(k + 1) * nc);
    }
    // printAvgMaxMin(mesh, field) of Utility.C:4-10 for moment k or its (scrubbed) source, as of the last
    // correct(): {avg, max, min} (global when the context carries a communicator)
    struct AvgMaxMin {
        double avg, max, min;
    };
    AvgMaxMin avgMaxMin(int k, bool source = false) const {
        const double *o = stats_.data() + (size_t)4 * ((source ? 0 : 2 * order_) + k);
        return {o[0] / o[1], o[2], o[3]};
    }
    AvgMaxMin avgMaxMinDiameter() const {
        const double *o = stats_.data() + (size_t)16 * order_;
        return {o[0] / o[1], o[2], o[3]};
    }

private:
    pbe_qmom *h_ = nullptr;
    pbe_qmom_state *state_ = nullptr;
    int order_ = 0;
    bool usingMULES_ = false;
    pbe_mesh *mesh_ = nullptr;
    int scheme_ = 0;
    std::vector<double> d_host_, stats_;
};

// ---- methodOfMoments (MOM.H:55-96, MOM.C:60-236): three moments, presumed gamma density -----------------
class MOM : public PBEMethod {
public:
    static constexpr const char *typeName = "methodOfMoments";

    // m0: initial moment fields m0, m1, m2, HOST [3][ncells] (MUST_READ "m<i>", MOM.C:128-141)
    MOM(const Dictionary &pbeProperties, const Phase &phase, const double *m0 = nullptr)
        : PBEMethod(pbeProperties, phase) {
        const Dictionary &c = pbeProperties.subDict("MOMCoeffs"); // MOM.C:68
        // the reference reads nothing from MOMCoeffs; quadraturePanels sets the fixed rule that
        // stands in for integrate()'s adaptive QAGIU (0 = PBE_MOM_DEFAULT_PANELS)
        const int panels = (int)c.lookupOrDefault<double>("quadraturePanels", 0.0);
        check(pbe_mom_create(ctx_.get(), panels, &spec_, &h_), "MOM::MOM");
        stageCellFields();
        const int64_t nc = phase.ncells;
        if (phase.mesh && nc) { // MOM::correct() ends in fvm::ddt + fvm::div(phi, m_i) == S_i for the 3 moments (MOM.C:211-222)
            if (!phase.V) throw FatalError("MOM: transport needs the cell volumes (Phase::V)");
            check(pbe_div_scheme_type(phase.mesh->implicitDivScheme.c_str(), &scheme_), "MOM::MOM div(phi,m.*)");
            mesh_ = makeDeviceMesh(ctx_.get(), phase, scheme_, "MOM::MOM mesh");
            const int64_t nf = phase.mesh->nInternalFaces + phase.mesh->nBoundaryFaces;
            mNew_ = DeviceArray(ctx_, 3 * nc), phi_ = DeviceArray(ctx_, nf);
            if (phase.mesh->boundaryValues) psib_ = DeviceArray(ctx_, 3 * phase.mesh->nBoundaryFaces);
        }
        m_ = DeviceArray(ctx_, 3 * nc), S_ = DeviceArray(ctx_, 3 * nc), gamma_ = DeviceArray(ctx_, 3 * nc);
        d_dev_ = DeviceArray(ctx_, nc);
        if (m0)
            m_.upload(m0, 3 * nc);
        else
            m_.zero();
        d_host_.assign((size_t)nc, 0.0);
        if (m0) { // d_ = pow(6/pi m1/m0, 1/3) at construction, MOM.C:143
            for (int64_t c2 = 0; c2 < nc; ++c2)
                d_host_[(size_t)c2] = std::pow(6.0 / M_PI * m0[nc + c2] / m0[c2], 1.0 / 3.0);
            d_stale_ = false;
        }
    }
    ~MOM() override {
        if (mesh_) pbe_mesh_destroy(mesh_);
        if (h_) pbe_mom_destroy(h_);
    }

    // MOM::correct(): gamma parameters, moment sources, NaN scrub, zero-flux update, clamp, d (MOM.C:148-236)
    void correct() override {
        const int64_t nc = phase_.ncells;
        if (nc == 0) return;
        if (mesh_) { // sources, NaN scrub (MOM.C:198-207), the implicit solve (:211-222), clamp (:224-231), d (:233)
            const FaceMesh &fm = *phase_.mesh;
            check(pbe_mom_source(h_, nc, nc, m_.data(), eps_.data(), rho_.data(), nu_.data(), S_.data(), gamma_.data(), nullptr),
                  "MOM::correct");
            check(pbe_field_scrub_nan(ctx_.get(), 3 * nc, S_.data(), nullptr), "MOM::correct");
            phi_.upload(fm.phi, fm.nInternalFaces + fm.nBoundaryFaces);
            if (fm.boundaryValues) psib_.upload(fm.boundaryValues, 3 * fm.nBoundaryFaces);
            const pbe_fvm_controls ctl = fvmControls(fm);
            perf_.resize(3);
            check(pbe_transport_implicit(mesh_, scheme_, 3, nc, m_.data(), phi_.data(), fm.boundaryValues ? psib_.data() : nullptr,
                                         fm.nBoundaryFaces, S_.data(), phase_.deltaT, &ctl, mNew_.data(), perf_.data(), nullptr),
                  "MOM::correct");
            std::swap(m_, mNew_);
            check(pbe_qmom_clamp_d(ctx_.get(), 3, nc, nc, m_.data(), d_dev_.data(), nullptr), "MOM::correct");
            d_stale_ = true;
            return;
        }
        check(pbe_mom_step(h_, nc, nc, m_.data(), eps_.data(), rho_.data(), nu_.data(), phase_.deltaT, S_.data(),
                           d_dev_.data(), gamma_.data(), nullptr),
              "MOM::correct");
        d_stale_ = true;
    }
    const std::vector<double> &d() const override {
        if (d_stale_ && phase_.ncells) {
            check(pbe_sync(ctx_.get()), "MOM::d");
            d_dev_.download(d_host_.data(), phase_.ncells);
            d_stale_ = false;
        }
        return d_host_;
    }
    // field k of: moments (m0, m1, m2), their scrubbed sources, or gamma_c0 / gamma_k / gamma_theta
    enum Field { Moment, Source, Gamma };
    std::vector<double> field(Field which, int k) const {
        std::vector<double> v((size_t)phase_.ncells);
        check(pbe_sync(ctx_.get()), "MOM::field");
        const DeviceArray &a = which == Moment ? m_ : which == Source ? S_ : gamma_;
        if (phase_.ncells) a.download(v.data(), phase_.ncells, (int64_t)k * phase_.ncells);
        return v;
    }

private:
    pbe_mom *h_ = nullptr;
    pbe_mesh *mesh_ = nullptr;
    int scheme_ = 0;
    std::vector<pbe_fvm_perf> perf_;
    DeviceArray m_, S_, gamma_, d_dev_, mNew_, phi_, psib_;
    mutable std::vector<double> d_host_;
    mutable bool d_stale_ = true;
};

// ---- dummy (dummy.H:55-79): constant diameter from dummyCoeffs.constantDiameter --------------------------
class dummy : public PBEMethod {
public:
    static constexpr const char *typeName = "dummy";
    dummy(const Dictionary &pbeProperties, const Phase &phase)
        : PBEMethod(pbeProperties, phase),
          d_((size_t)phase.ncells, pbeProperties.subDict("dummyCoeffs").readScalar("constantDiameter")) {}
    void correct() override {}
    const std::vector<double> &d() const override { return d_; }

private:
    std::vector<double> d_;
};

} // namespace PBEMethods

// addToRunTimeSelectionTable(PBEMethod, X, dictionary): MOC.C:38-39, QMOM.C:50-51, MOM.C:50-51, dummy.C:37-38
inline void registerBuiltinMethods() {
    auto &t = PBEMethod::table();
    if (!t.empty()) return;
    t[PBEMethods::MOC::typeName] = [](const Dictionary &d, const Phase &p) {
        return std::unique_ptr<PBEMethod>(new PBEMethods::MOC(d, p));
    };
    t[PBEMethods::QMOM::typeName] = [](const Dictionary &d, const Phase &p) {
        return std::unique_ptr<PBEMethod>(new PBEMethods::QMOM(d, p));
    };
    t[PBEMethods::MOM::typeName] = [](const Dictionary &d, const Phase &p) {
        return std::unique_ptr<PBEMethod>(new PBEMethods::MOM(d, p));
    };
    t[PBEMethods::dummy::typeName] = [](const Dictionary &d, const Phase &p) {
        return std::unique_ptr<PBEMethod>(new PBEMethods::dummy(d, p));
    };
}

// PBEMethod::New (PBEMethod.C:81-108): word(dict.lookup("method")) -> table lookup; unknown names
// are fatal and list the valid ones.
inline std::unique_ptr<PBEMethod> PBEMethod::New(const Dictionary &pbeProperties, const Phase &phase) {
    registerBuiltinMethods();
    const std::string method = pbeProperties.lookup("method");
    auto it = table().find(method);
    if (it == table().end()) {
        std::ostringstream os;
        os << "--> PBE FATAL ERROR in PBEMethod::New : Unknown pbeModelType type " << method
           << "\n\nValid PBEMethod types are :\n(";
        bool first = true;
        for (auto &kv : table()) {
            os << (first ? "" : " ") << kv.first;
            first = false;
        }
        os << ")";
        throw FatalError(os.str());
    }
    return it->second(pbeProperties, phase);
}

} // namespace pbe
