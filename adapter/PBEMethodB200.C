// PBEMethodB200.C -- see PBEMethodB200.H.  Every block cites the reference lines it replaces.
#include "PBEMethodB200.H"

#include <cstring>

namespace Foam {

defineTypeNameAndDebug(PBEMethod, 0);
defineRunTimeSelectionTable(PBEMethod, dictionary);

void b200Check(int rc, const char *where) {
    if (rc != PBE_OK) FatalErrorIn(where) << pbe_last_error() << exit(FatalError);
}

// coalescenceKernel::New / breakupKernel::New / daughterParticleDistribution::New
// (coalescenceKernel.C:52-70, breakupKernel.C:52-70, daughterParticleDistribution.C:63-81): the
// TypeName is looked up in the library's tables -- an unknown one fails listing the valid types -- and
// the coefficients come from pbeDict.subDict(<TypeName> + "Coeffs"), which is fatal when absent.
pbe_kernel_spec b200KernelSpec(const dictionary &pbeProperties) {
    pbe_kernel_spec s;
    std::memset(&s, 0, sizeof(s));
    const word ck(pbeProperties.lookup("coalescenceKernel"));
    Info << "Selecting coalescence model: " << ck << endl;
    b200Check(pbe_coalescence_type(ck.c_str(), &s.coalescence), "coalescenceKernel::New");
    const dictionary &cd = pbeProperties.subDict(ck + "Coeffs");
    if (s.coalescence == PBE_COAL_CONSTANT) {
        s.coal_c[0] = readScalar(cd.lookup("constant")); // constantCoalescence.C:48
    } else if (s.coalescence == PBE_COAL_CT) {           // coalescenceKernels/CoulaloglouTavlarides.C:39-42
        s.coal_c[0] = cd.lookupOrDefault<scalar>("c1", 0.00487);
        s.coal_c[1] = cd.lookupOrDefault<scalar>("c2", 0.008);
        s.coal_c[2] = cd.lookupOrDefault<scalar>("gamma", 0.0);
        s.coal_c[3] = cd.lookupOrDefault<scalar>("sigma", 0.047);
    }
    const word bk(pbeProperties.lookup("breakupKernel"));
    Info << "Selecting breakup model: " << bk << endl;
    b200Check(pbe_breakup_type(bk.c_str(), &s.breakup), "breakupKernel::New");
    const dictionary &bd = pbeProperties.subDict(bk + "Coeffs");
    if (s.breakup == PBE_BR_CT) { // breakupKernels/CoulaloglouTavlarides.C:46-49
        s.br_c[0] = bd.lookupOrDefault<scalar>("c1", 0.00487);
        s.br_c[1] = bd.lookupOrDefault<scalar>("c2", 0.008);
        s.br_c[2] = bd.lookupOrDefault<scalar>("gamma", 0.0);
        s.br_c[3] = bd.lookupOrDefault<scalar>("sigma", 0.047);
    }
    const word dk(pbeProperties.lookup("daughterParticleDistribution"));
    Info << "Selecting daughter particle model: " << dk << endl;
    b200Check(pbe_daughter_type(dk.c_str(), &s.daughter), "daughterParticleDistribution::New");
    (void)pbeProperties.subDict(dk + "Coeffs");
    return s;
}

// ---- PBEMethod (PBEMethod.C:40-108) ---------------------------------------------------------------
PBEMethod::PBEMethod(const dictionary &pbeProperties, const phaseModel &phase)
    : phase_(phase), spec_(b200KernelSpec(pbeProperties)), ctx_(nullptr), rhoHost_(phase.size()),
      nuHost_(phase.size()) {
    // one context per rank; with several GPUs per node a real build passes
    // Pstream::myProcNo() % nDevices
    b200Check(pbe_ctx_create(pbeProperties.lookupOrDefault<label>("device", 0), &ctx_), "PBEMethod::PBEMethod");
}

PBEMethod::~PBEMethod() {
    if (ctx_) pbe_ctx_destroy(ctx_);
}

autoPtr<PBEMethod> PBEMethod::New(const dictionary &pbeProperties, const phaseModel &phase) {
    word pbeMethod(pbeProperties.lookup("method"));
    Info << "Selecting population balance model: " << pbeMethod << endl;
    dictionaryConstructorTable *table = dictionaryConstructorTablePtr_();
    dictionaryConstructorTable::iterator cstrIter;
    if (!table || (cstrIter = table->find(pbeMethod)) == table->end()) {
        FatalErrorIn("PBEMethod::New")
            << "Unknown pbeModelType type " << pbeMethod << endl
            << endl
            << "Valid PBEMethod types are : " << endl
            << sortedToc(table) << exit(FatalError);
    }
    return cstrIter->second(pbeProperties, phase);
}

const scalar *PBEMethod::epsilonData() const {
    // coalescenceKernels/CoulaloglouTavlarides.C:52-53: epsilon.<phase> from the mesh registry
    return phase_.mesh().lookupObject<volScalarField>("epsilon." + phase_.name()).internalField().cdata();
}

void PBEMethod::updatePhaseFields() {
    // the reference's update() hooks (coalescenceKernels/CoulaloglouTavlarides.C:46): rho and nu
    // are tmp<> fields in twoPhaseEulerFoam, so they are evaluated once per time step here
    const tmp<volScalarField> trho(phase_.rho()), tnu(phase_.nu());
    std::memcpy(rhoHost_.data_ptr(), trho().internalField().cdata(), sizeof(scalar) * phase_.size());
    std::memcpy(nuHost_.data_ptr(), tnu().internalField().cdata(), sizeof(scalar) * phase_.size());
}

int PBEMethod::divSchemeOf(const dictionary &coeffs, const char *where) const {
#ifdef PBE_B200_REAL_OPENFOAM
    // fvSchemes: divSchemes { "div\\(phi,n.*\\)" Gauss vanLeer; }  (c01/system/fvSchemes:34)
    ITstream &is = phase_.mesh().divScheme("div(phi,n0)");
    word scheme(is);
    if (!is.eof()) scheme += " " + word(is);
#else
    const word scheme = phase_.mesh().divScheme_.empty() ? coeffs.lookupOrDefault<word>("divScheme", "Gauss upwind")
                                                         : phase_.mesh().divScheme_;
#endif
    int id = PBE_DIV_UPWIND;
    b200Check(pbe_div_scheme_type(scheme.c_str(), &id), where); // unknown: FatalError listing the valid ones
    return id;
}

// usingMULES false: the scheme of the implicit convection term -- fvSchemes "div\\(U,nk\\)" (MOC.C:194) or the default
// name fvm::div builds, "div(phi,m<k>)" (QMOM.C:203-204) -- and what fvSolution says about the equation: solvers {
// "n.*" { tolerance; relTol; minIter; maxIter; } } and relaxationFactors { equations { ".*" f; } } (what eqn.relax()
// looks up).  The linear solver named there (PBiCG + DILU, smoothSolver) is replaced by the device iteration; the
// stopping criterion is kept.  Returns false when the scheme has no device path (pbe_last_error() lists the valid ones).
bool PBEMethod::implicitControlsOf(const dictionary &coeffs, const word &field, const word &divName, int &scheme,
                                   pbe_fvm_controls &ctl) const {
    ctl.tolerance = 1e-6, ctl.rel_tol = 0.0, ctl.min_iter = 0, ctl.max_iter = 1000; // lduMatrix::solver defaults
    ctl.relaxation = 0.0, ctl.solver = PBE_FVM_SOLVER_AUTO;
#ifdef PBE_B200_REAL_OPENFOAM
    ITstream &is = phase_.mesh().divScheme(divName);
    word name(is);
    if (!is.eof()) name += " " + word(is);
    const dictionary &sd = phase_.mesh().solverDict(field);
    ctl.tolerance = sd.lookupOrDefault<scalar>("tolerance", 1e-6);
    ctl.rel_tol = sd.lookupOrDefault<scalar>("relTol", 0.0);
    ctl.min_iter = sd.lookupOrDefault<label>("minIter", 0);
    ctl.max_iter = sd.lookupOrDefault<label>("maxIter", 1000);
    if (phase_.mesh().relaxEquation(field)) ctl.relaxation = phase_.mesh().equationRelaxationFactor(field);
#else
    (void)field, (void)divName;
    const word name = coeffs.lookupOrDefault<word>("implicitDivScheme", "Gauss upwind");
    ctl.tolerance = coeffs.lookupOrDefault<scalar>("tolerance", 1e-6);
    ctl.rel_tol = coeffs.lookupOrDefault<scalar>("relTol", 0.0);
    ctl.min_iter = coeffs.lookupOrDefault<label>("minIter", 0);
    ctl.max_iter = coeffs.lookupOrDefault<label>("maxIter", 1000);
    ctl.relaxation = coeffs.lookupOrDefault<scalar>("relaxationFactor", 0.0);
#endif
    return pbe_div_scheme_type(name.c_str(), &scheme) == PBE_OK; // upwind, vanLeer, limitedLinear 1
}

pbe_mesh *PBEMethod::newDeviceMesh(int scheme, const char *where) const {
    const fvMesh &mesh = phase_.mesh();
    pbe_mesh *m = nullptr;
#ifdef PBE_B200_REAL_OPENFOAM
    std::vector<int32_t> own(mesh.owner().begin(), mesh.owner().end()), nei(mesh.neighbour().begin(), mesh.neighbour().end()), bfc;
    forAll(mesh.boundary(), patchi) {
        const labelUList &fc = mesh.boundary()[patchi].faceCells();
        bfc.insert(bfc.end(), fc.begin(), fc.end());
    }
    b200Check(pbe_mesh_create(ctx_, mesh.nCells(), mesh.nInternalFaces(), own.data(), nei.data(), (int64_t)bfc.size(),
                              bfc.data(), mesh.V().field().cdata(), &m), where);
    if (scheme != PBE_DIV_UPWIND) {
        std::vector<double> Sf(3 * mesh.nFaces()), C(3 * mesh.nCells());
        forAll(mesh.Sf().internalField(), f) for (int d = 0; d < 3; ++d) Sf[3 * f + d] = mesh.Sf().internalField()[f][d];
        label f0 = mesh.nInternalFaces();
        forAll(mesh.boundary(), patchi) forAll(mesh.boundary()[patchi], i) {
            for (int d = 0; d < 3; ++d) Sf[3 * f0 + d] = mesh.Sf().boundaryField()[patchi][i][d];
            ++f0;
        }
        forAll(mesh.C().internalField(), c) for (int d = 0; d < 3; ++d) C[3 * c + d] = mesh.C().internalField()[c][d];
        b200Check(pbe_mesh_set_geometry(m, Sf.data(), C.data(), mesh.weights().internalField().cdata()), where);
    }
#else
    std::vector<int32_t> own(mesh.owner().begin(), mesh.owner().end()), nei(mesh.neighbour().begin(), mesh.neighbour().end()),
        bfc(mesh.boundaryFaceCells_.begin(), mesh.boundaryFaceCells_.end());
    b200Check(pbe_mesh_create(ctx_, mesh.nCells(), mesh.nInternalFaces(), own.data(), nei.data(), (int64_t)bfc.size(),
                              bfc.data(), mesh.V().cdata(), &m), where);
    if (scheme != PBE_DIV_UPWIND) {
        if (!mesh.hasGeometry())
            FatalErrorIn(where) << "the limited divergence scheme needs fvMesh::Sf(), C() and weights()" << exit(FatalError);
        b200Check(pbe_mesh_set_geometry(m, mesh.Sf_.data(), mesh.C_.data(), mesh.weights_.data()), where);
    }
#endif
    return m;
}

namespace PBEMethods {

static inline std::string fieldName(const char *stem, label i) {
    std::ostringstream os;
    os << stem << i;
    return os.str();
}
//- printAvgMaxMin (Utility.C:4-10) from the {sum f V, sum V, max, min} partials the library returns
//  (a parallel run reduces the four numbers over the ranks first: sum, sum, max, min)
static inline void printAvgMaxMin(const word &name, const scalar *s) {
    Info << name << ": avg, max,min " << s[0] / s[1] << ", " << s[2] << ", " << s[3] << endl;
}

// ---- MOC ------------------------------------------------------------------------------------------
defineTypeNameAndDebug(MOC, 0);
addToRunTimeSelectionTable(PBEMethod, MOC, dictionary);

MOC::MOC(const dictionary &pbeProperties, const phaseModel &phase)
    : PBEMethod(pbeProperties, phase), runTime(phase.mesh().time()), deltaXi_("deltaXi", 0.0),
      numberOfClasses_(0), usingMULES_(false), source_term_relaxation_(1.0), moc_(nullptr), state_(nullptr),
      b200Mesh_(nullptr), implicitOnDevice_(false), d_("d." + phase.name(), phase.mesh(), 0.0, false), continuityError_(0.0), hostStale_(false),
      deviceStale_(true) {
    const dictionary &MOCDict_ = pbeProperties.subDict("MOCCoeffs"); // MOC.C:51-60
    numberOfClasses_ = readLabel(MOCDict_.lookup("numberOfClasses"));
    deltaXi_ = dimensionedScalar("deltaXi", readScalar(MOCDict_.lookup("xi1")));
    usingMULES_ = MOCDict_.lookupOrDefault<bool>("usingMULES", false);
    source_term_relaxation_ = MOCDict_.lookupOrDefault<scalar>("sourceRelaxation", 1.0);
    Info << "Creating " << numberOfClasses_ << " class" << (numberOfClasses_ > 1 ? "es" : "") << endl;
    classNumberDensity_ = PtrList<volScalarField>(numberOfClasses_);
    classMin_ = scalarField(numberOfClasses_, 0.0);
    forAll(classNumberDensity_, i) // MOC.C:69-77: fields n0 .. n<N-1>, MUST_READ, registered on the mesh
        classNumberDensity_.set(i, new volScalarField(fieldName("n", i), phase.mesh()));
    // EXTENSION (no counterpart in the reference's MOCCoeffs): classGridRatio r > 1 -> geometric class
    // volumes xi1 r^i through the fixed-pivot technique; absent: the reference's uniform grid (i + 1) xi1 (MOC.C:56,79)
    const scalar gridRatio = MOCDict_.lookupOrDefault<scalar>("classGridRatio", 0.0);
    if (gridRatio > 0.0) {
        if (!(gridRatio > 1.0))
            FatalErrorIn("MOC::MOC") << "classGridRatio must be > 1 (got " << gridRatio << ")" << exit(FatalError);
        std::vector<double> piv((size_t)numberOfClasses_);
        forAll(classNumberDensity_, i) piv[(size_t)i] = (i == 0) ? deltaXi_.value() : piv[(size_t)i - 1] * gridRatio;
        b200Check(pbe_moc_create_pivots(ctx_, numberOfClasses_, piv.data(), &spec_, source_term_relaxation_, &moc_),
                  "MOC::MOC");
    } else {
        b200Check(pbe_moc_create(ctx_, numberOfClasses_, deltaXi_.value(), &spec_, source_term_relaxation_, &moc_),
                  "MOC::MOC");
    }
    const fvMesh &mesh = phase.mesh();
    b200Check(pbe_moc_state_create(moc_, mesh.nCells(), mesh.V().cdata(), &state_), "MOC::MOC");
    if (mesh.hasFaces()) {
        if (!usingMULES_) { // solveWithFVM (MOC.C:177-180, 186-198), the reference's default
            int scheme = PBE_DIV_UPWIND;
            pbe_fvm_controls ctl;
            if (implicitControlsOf(MOCDict_, fieldName("n", 0), "div(U,nk)", scheme, ctl)) {
                b200Mesh_ = newDeviceMesh(scheme, "MOC::MOC mesh");
                b200Check(pbe_moc_state_attach_mesh(state_, b200Mesh_, scheme), "MOC::MOC");
                b200Check(pbe_moc_state_set_implicit(state_, &ctl), "MOC::MOC usingMULES false");
                implicitOnDevice_ = true;
            } else {
#ifdef PBE_B200_REAL_OPENFOAM
                // a scheme without a device path: the solve stays on the host, see correct()
                Info << "MOC: div(U,nk) has no device path: sources on the device, fvm solve on the host" << endl;
#else
                FatalErrorIn("MOC::MOC") << "usingMULES false, div(U,nk) of the implicit solve (MOC.C:186-198): "
                                         << pbe_last_error() << exit(FatalError);
#endif
            }
        } else {
            const int scheme = divSchemeOf(MOCDict_, "MOC::MOC div(phi,n.*)");
            b200Mesh_ = newDeviceMesh(scheme, "MOC::MOC mesh");
            b200Check(pbe_moc_state_attach_mesh(state_, b200Mesh_, scheme), "MOC::MOC");
        }
    } else if (!usingMULES_) {
        // no faces: fvm::ddt(n) == S and MULES::explicitSolve both reduce to n += dt S
        Info << "MOC: no face fluxes on this mesh: explicit zero-flux update (usingMULES false would solve the same "
                "diagonal system)" << endl;
    }
}

MOC::~MOC() {
    if (state_) pbe_moc_state_destroy(state_);
    if (b200Mesh_) pbe_mesh_destroy(b200Mesh_);
    if (moc_) pbe_moc_destroy(moc_);
}

void MOC::upload() { // each field is contiguous: it becomes row k of the class-major device block
    const label nCells = phase_.size();
    std::vector<double> stage((size_t)numberOfClasses_ * nCells);
    forAll(classNumberDensity_, k)
        std::memcpy(stage.data() + (size_t)k * nCells, classNumberDensity_[k].internalField().cdata(), sizeof(double) * nCells);
    b200Check(pbe_moc_state_upload(state_, nCells, stage.data()), "MOC::upload");
    deviceStale_ = false;
    hostStale_ = false;
}

void MOC::download() const { // runTime.write(): the AUTO_WRITE fields n<k>
    const label nCells = phase_.size();
    std::vector<double> stage((size_t)numberOfClasses_ * nCells);
    b200Check(pbe_moc_state_download(state_, nCells, stage.data()), "MOC::download");
    forAll(classNumberDensity_, k)
        std::memcpy(classNumberDensity_[k].internalField().data_ptr(), stage.data() + (size_t)k * nCells, sizeof(double) * nCells);
    hostStale_ = false;
}

volScalarField &MOC::classNumberDensity(label k) {
    if (hostStale_) download();
    deviceStale_ = true; // the caller may write it
    return classNumberDensity_[k];
}

// MOC::correct(), MOC.C:159-184, in one library call: the breakup-rate cache fill (:162-171),
// coalescence_().update() (:173), the N classSourceTerm(k) (:174-175, i.e. :87-157), solveWithMULES (:198-272: limited
// high-order flux, MULES::limit, limitSum, explicitSolve -- or the zero-flux update when the mesh has no
// faces), the continuity check (:254-279), the min(n_k) prints (:181-183) and d() (:282-290).
void MOC::correct() {
    if (deviceStale_) upload();
    updatePhaseFields();
#ifdef PBE_B200_REAL_OPENFOAM
    if (!usingMULES_ && !implicitOnDevice_ && phase_.mesh().nInternalFaces() > 0) {
        // implicit path with a limited scheme: the sources come from the device, the reference's solveWithFVM(S) (MOC.C:186-198)
        // runs unchanged on the host fields, which are then uploaded again
        const label nCells = phase_.size();
        std::vector<double> nH((size_t)numberOfClasses_ * nCells), SH(nH.size());
        b200Check(pbe_moc_state_download(state_, nCells, nH.data()), "MOC::correct");
        b200Check(pbe_moc_source_host(moc_, nCells, nCells, nH.data(), epsilonData(), rhoHost_.cdata(), nuHost_.cdata(),
                                      runTime.deltaT().value(), SH.data()), "MOC::correct");
        download();
        PtrList<volScalarField> S(numberOfClasses_);
        forAll(classNumberDensity_, k) {
            S.set(k, new volScalarField(IOobject("S", runTime.timeName(), phase_.mesh()), phase_.mesh(),
                                        dimensionedScalar("S", classNumberDensity_[k].dimensions() / dimTime, 0)));
            std::memcpy(S[k].internalField().begin(), SH.data() + (size_t)k * nCells, sizeof(double) * nCells);
        }
        solveWithFVM(S);
        deviceStale_ = true;
        d_ = d();
        return;
    }
#endif
    pbe_moc_diag diag;
    std::memset(&diag, 0, sizeof(diag));
    diag.class_min = classMin_.data_ptr();
    b200Check(pbe_moc_correct_resident(state_, epsilonData(), rhoHost_.cdata(), nuHost_.cdata(),
                                       phase_.internalField().cdata(), b200Mesh_ ? phase_.phi().cdata() : nullptr,
                                       nullptr /* zeroGradient patches */, 0, runTime.deltaT().value(),
                                       d_.internalField().data_ptr(), &diag),
              "MOC::correct");
    hostStale_ = true;
    continuityError_ = diag.cont_err_max;
    Info << "Max continuity error between advection and PBE: " << continuityError_ << endl; // MOC.C:278-279
    forAll(classNumberDensity_, k) // MOC.C:181-183 (a parallel run takes the min over the ranks)
        Info << "min(n" << k << ") " << classMin_[k] << endl;
    if (runTime.outputTime()) download(); // the solver writes the registered fields after this step
}

// MOC::d(), MOC.C:282-290 with volumeToDiameter (:43-45): number-mean diameter of the CURRENT fields
const volScalarField MOC::d() const {
    if (deviceStale_ || !hostStale_) { // nothing stepped since the fields were last written: recompute from them
        MOC &self = const_cast<MOC &>(*this);
        if (deviceStale_) self.upload();
        b200Check(pbe_moc_state_d(state_, self.d_.internalField().data_ptr()), "MOC::d");
    }
    return d_;
}

// ---- QMOM -----------------------------------------------------------------------------------------
defineTypeNameAndDebug(QMOM, 0);
addToRunTimeSelectionTable(PBEMethod, QMOM, dictionary);

QMOM::QMOM(const dictionary &pbeProperties, const phaseModel &phase)
    : PBEMethod(pbeProperties, phase), runTime(phase.mesh().time()),
      quadratureOrder_(readLabel(pbeProperties.subDict("QMOMCoeffs").lookup("quadratureOrder"))), // QMOM.C:63
      usingMULES_(pbeProperties.subDict("QMOMCoeffs").lookupOrDefault<bool>("usingMULES", false)),
      moments_(2 * quadratureOrder_), qmom_(nullptr), state_(nullptr), b200Mesh_(nullptr),
      d_("diameter", phase.mesh(), 0.0), stats_((4 * quadratureOrder_ + 1) * 4, 0.0), hostStale_(false), deviceStale_(true) {
    forAll(moments_, k) moments_.set(k, new volScalarField(fieldName("m", k), phase.mesh())); // QMOM.C:72-79
    b200Check(pbe_qmom_create(ctx_, quadratureOrder_, &spec_, &qmom_), "QMOM::QMOM");
    const fvMesh &mesh = phase.mesh();
    b200Check(pbe_qmom_state_create(qmom_, mesh.nCells(), mesh.V().cdata(), &state_), "QMOM::QMOM");
    if (mesh.hasFaces()) {
        int scheme = PBE_DIV_UPWIND;
        pbe_fvm_controls ctl;
        const dictionary &coeffs = pbeProperties.subDict("QMOMCoeffs");
        if (!usingMULES_) { // solveWithFVM (QMOM.C:112-115, 201-209), the reference's default
#ifdef PBE_B200_REAL_OPENFOAM
            const word divName = "div(" + phase.phi().name() + ',' + fieldName("m", 0) + ')'; // fvm::div's default name
#else
            const word divName = "div(phi,m0)";
#endif
            if (!implicitControlsOf(coeffs, fieldName("m", 0), divName, scheme, ctl))
                FatalErrorIn("QMOM::QMOM") << "usingMULES false, div(phi,m.*) of the implicit solve (QMOM.C:201-209): "
                                           << pbe_last_error() << exit(FatalError);
        } else {
            scheme = divSchemeOf(coeffs, "QMOM::QMOM div(phi,m.*)");
        }
        b200Mesh_ = newDeviceMesh(scheme, "QMOM::QMOM mesh");
        b200Check(pbe_qmom_state_attach_mesh(state_, b200Mesh_, scheme), "QMOM::QMOM");
        if (!usingMULES_) b200Check(pbe_qmom_state_set_implicit(state_, &ctl), "QMOM::QMOM usingMULES false");
    }
}

QMOM::~QMOM() {
    if (state_) pbe_qmom_state_destroy(state_);
    if (b200Mesh_) pbe_mesh_destroy(b200Mesh_);
    if (qmom_) pbe_qmom_destroy(qmom_);
}

void QMOM::upload() {
    const label nCells = phase_.size();
    std::vector<double> stage((size_t)moments_.size() * nCells);
    forAll(moments_, k)
        std::memcpy(stage.data() + (size_t)k * nCells, moments_[k].internalField().cdata(), sizeof(double) * nCells);
    b200Check(pbe_qmom_state_upload(state_, nCells, stage.data()), "QMOM::upload");
    b200Check(pbe_qmom_state_d(state_, d_.internalField().data_ptr()), "QMOM::upload"); // d_ = pow(6/pi m1/m0, 1/3), QMOM.C:81
    deviceStale_ = false;
    hostStale_ = false;
}

void QMOM::download() const {
    const label nCells = phase_.size();
    std::vector<double> stage((size_t)moments_.size() * nCells);
    b200Check(pbe_qmom_state_download(state_, nCells, stage.data()), "QMOM::download");
    forAll(moments_, k)
        std::memcpy(moments_[k].internalField().data_ptr(), stage.data() + (size_t)k * nCells, sizeof(double) * nCells);
    hostStale_ = false;
}

volScalarField &QMOM::moment(label k) {
    if (hostStale_) download();
    deviceStale_ = true;
    return moments_[k];
}

// QMOM::correct(), QMOM.C:85-129, in one library call: the 2N momentSourceTerm(k) (:88-96, i.e. :215-267 with one
// inversion per cell instead of one per moment index), the NaN scrub (:98-107), solveWithMULES (:131-197) or the
// zero-flux update, the SMALL clamp (:117-123), d_ (:125) and the printAvgMaxMin partials (:109, :122, :128).
void QMOM::correct() {
    if (deviceStale_) upload();
    updatePhaseFields();
    const label nm = moments_.size();
    pbe_qmom_diag diag;
    diag.source = stats_.data_ptr();
    diag.moment = stats_.data_ptr() + 4 * nm;
    diag.d = stats_.data_ptr() + 8 * nm;
    b200Check(pbe_qmom_correct_resident(state_, epsilonData(), rhoHost_.cdata(), nuHost_.cdata(),
                                        b200Mesh_ ? phase_.phi().cdata() : nullptr, nullptr, 0, runTime.deltaT().value(),
                                        d_.internalField().data_ptr(), &diag),
              "QMOM::correct");
    hostStale_ = true;
    Info << "moment sources:" << endl; // QMOM.C:97-110
    for (label k = 0; k < nm; ++k) printAvgMaxMin(fieldName("m", k) + "Source", diag.source + 4 * k);
    for (label k = 0; k < nm; ++k) printAvgMaxMin(fieldName("m", k), diag.moment + 4 * k); // QMOM.C:117-123
    printAvgMaxMin("diameter", diag.d);                                                      // QMOM.C:128
    if (runTime.outputTime()) download();
}

// ---- MOM ------------------------------------------------------------------------------------------
defineTypeNameAndDebug(MOM, 0);
addToRunTimeSelectionTable(PBEMethod, MOM, dictionary);

MOM::MOM(const dictionary &pbeProperties, const phaseModel &phase)
    : PBEMethod(pbeProperties, phase), runTime(phase.mesh().time()), moments_(3), mom_(nullptr), mDev_(nullptr),
      epsDev_(nullptr), rhoDev_(nullptr), nuDev_(nullptr), dDev_(nullptr), SDev_(nullptr), mNewDev_(nullptr), phiDev_(nullptr),
      b200Mesh_(nullptr), scheme_(PBE_DIV_UPWIND), d_("diameter", phase.mesh(), 0.0) {
    const dictionary &MOMDict_ = pbeProperties.subDict("MOMCoeffs"); // MOM.C:68
    forAll(moments_, k) moments_.set(k, new volScalarField(fieldName("m", k), phase.mesh())); // MOM.C:128-141
    // the reference reads nothing from MOMCoeffs; quadraturePanels sizes the fixed rule standing in for QAGIU
    b200Check(pbe_mom_create(ctx_, MOMDict_.lookupOrDefault<label>("quadraturePanels", 0), &spec_, &mom_), "MOM::MOM");
    const int64_t row = (int64_t)sizeof(double) * (phase.size() > 0 ? phase.size() : 1);
    b200Check(pbe_device_alloc(ctx_, 3 * row, (void **)&mDev_), "MOM::MOM");
    if (phase.mesh().hasFaces()) { // MOM.C:211-222: fvm::ddt + fvm::div(phi, m_i) == S_i, relax(), solve() for the 3 moments
#ifdef PBE_B200_REAL_OPENFOAM
        const word divName = "div(" + phase.phi().name() + ',' + fieldName("m", 0) + ')';
#else
        const word divName = "div(phi,m0)";
#endif
        if (!implicitControlsOf(MOMDict_, fieldName("m", 0), divName, scheme_, fvm_))
            FatalErrorIn("MOM::MOM") << "div(phi,m.*) of the implicit solve (MOM.C:211-222): " << pbe_last_error()
                                     << exit(FatalError);
        b200Mesh_ = newDeviceMesh(scheme_, "MOM::MOM mesh");
        b200Check(pbe_device_alloc(ctx_, 3 * row, (void **)&SDev_), "MOM::MOM");
        b200Check(pbe_device_alloc(ctx_, 3 * row, (void **)&mNewDev_), "MOM::MOM");
        b200Check(pbe_device_alloc(ctx_, (int64_t)sizeof(double) * phase.phi().size(), (void **)&phiDev_), "MOM::MOM");
    }
    for (double **p : {&epsDev_, &rhoDev_, &nuDev_, &dDev_}) b200Check(pbe_device_alloc(ctx_, row, (void **)p), "MOM::MOM");
}

MOM::~MOM() {
    for (double *p : {mDev_, epsDev_, rhoDev_, nuDev_, dDev_, SDev_, mNewDev_, phiDev_})
        if (p) pbe_device_free(ctx_, p);
    if (b200Mesh_) pbe_mesh_destroy(b200Mesh_);
    if (mom_) pbe_mom_destroy(mom_);
}

// MOM::correct(), MOM.C:148-236: gamma parameters, the nested-integral moment sources, NaN scrub, zero-flux
// update, SMALL clamp and d on the device (pbe_mom_step); the three moment fields travel both ways every step
void MOM::correct() {
    Info << "updating size moments" << endl; // MOM.C:150
    const label nCells = phase_.size();
    if (nCells == 0) return;
    updatePhaseFields();
    const int64_t row = (int64_t)sizeof(double) * nCells;
    forAll(moments_, k)
        b200Check(pbe_memcpy_h2d(ctx_, mDev_ + (size_t)k * nCells, moments_[k].internalField().cdata(), row), "MOM::correct");
    b200Check(pbe_memcpy_h2d(ctx_, epsDev_, epsilonData(), row), "MOM::correct");
    b200Check(pbe_memcpy_h2d(ctx_, rhoDev_, rhoHost_.cdata(), row), "MOM::correct");
    b200Check(pbe_memcpy_h2d(ctx_, nuDev_, nuHost_.cdata(), row), "MOM::correct");
    if (b200Mesh_) { // sources, NaN scrub (MOM.C:198-207), the implicit solve (:211-222), clamp (:224-231), d (:233)
        b200Check(pbe_mom_source(mom_, nCells, nCells, mDev_, epsDev_, rhoDev_, nuDev_, SDev_, nullptr, nullptr), "MOM::correct");
        b200Check(pbe_field_scrub_nan(ctx_, (int64_t)3 * nCells, SDev_, nullptr), "MOM::correct");
        b200Check(pbe_memcpy_h2d(ctx_, phiDev_, phase_.phi().cdata(), (int64_t)sizeof(double) * phase_.phi().size()), "MOM::correct");
        b200Check(pbe_transport_implicit(b200Mesh_, scheme_, 3, nCells, mDev_, phiDev_, nullptr /* zeroGradient patches */, 0,
                                         SDev_, runTime.deltaT().value(), &fvm_, mNewDev_, nullptr, nullptr),
                  "MOM::correct");
        std::swap(mDev_, mNewDev_);
        b200Check(pbe_qmom_clamp_d(ctx_, 3, nCells, nCells, mDev_, dDev_, nullptr), "MOM::correct");
    } else {
        b200Check(pbe_mom_step(mom_, nCells, nCells, mDev_, epsDev_, rhoDev_, nuDev_, runTime.deltaT().value(), nullptr, dDev_,
                               nullptr, nullptr),
                  "MOM::correct");
    }
    b200Check(pbe_sync(ctx_), "MOM::correct");
    forAll(moments_, k)
        b200Check(pbe_memcpy_d2h(ctx_, moments_[k].internalField().data_ptr(), mDev_ + (size_t)k * nCells, row), "MOM::correct");
    b200Check(pbe_memcpy_d2h(ctx_, d_.internalField().data_ptr(), dDev_, row), "MOM::correct");
}

// ---- dummy (dummy.C:45-70) ------------------------------------------------------------------------
defineTypeNameAndDebug(dummy, 0);
addToRunTimeSelectionTable(PBEMethod, dummy, dictionary);

dummy::dummy(const dictionary &pbeProperties, const phaseModel &phase)
    : PBEMethod(pbeProperties, phase),
      d_("d." + phase.name(), phase.mesh(), readScalar(pbeProperties.subDict("dummyCoeffs").lookup("constantDiameter")), false) {}

} // namespace PBEMethods

// ---- PBEDiameter (PBEDiameter.C:36-81, PBEDiameter.H:95-112) ------------------------------------------
namespace diameterModels {
defineTypeNameAndDebug(PBEDiameter, 0);
addToRunTimeSelectionTable(diameterModel, PBEDiameter, dictionary);

PBEDiameter::PBEDiameter(const dictionary &diameterProperties, const phaseModel &phase)
    : diameterModel(diameterProperties, phase), method_(PBEMethod::New(diameterProperties, phase)),
      d_("d." + phase.name(), phase.mesh(), 0.0) {
    // d_(IOobject(...), method_->d()), PBEDiameter.C:64-75
    const volScalarField d0(method_->d());
    forAll(d_.internalField(), c) d_[c] = d0[c];
}

void PBEDiameter::correct() {
    method_->correct();
    const volScalarField d(method_->d());
    forAll(d_.internalField(), c) d_[c] = d[c];
}
} // namespace diameterModels

} // namespace Foam
