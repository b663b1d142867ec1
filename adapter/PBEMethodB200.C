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
// TypeName is looked up in the library's tables; an unknown one fails listing the valid types.
pbe_kernel_spec b200KernelSpec(const dictionary &pbeProperties) {
    pbe_kernel_spec s;
    std::memset(&s, 0, sizeof(s));
    const word ck(pbeProperties.lookup("coalescenceKernel"));
    b200Check(pbe_coalescence_type(ck.c_str(), &s.coalescence), "coalescenceKernel::New");
    const dictionary &cd = pbeProperties.optionalSubDict(ck + "Coeffs");
    if (s.coalescence == PBE_COAL_CONSTANT) {
        s.coal_c[0] = readScalar(cd.lookup("constant")); // constantCoalescence.C:48
    } else if (s.coalescence == PBE_COAL_CT) {           // coalescenceKernels/CoulaloglouTavlarides.C:39-42
        s.coal_c[0] = cd.lookupOrDefault<scalar>("c1", 0.00487);
        s.coal_c[1] = cd.lookupOrDefault<scalar>("c2", 0.008);
        s.coal_c[2] = cd.lookupOrDefault<scalar>("gamma", 0.0);
        s.coal_c[3] = cd.lookupOrDefault<scalar>("sigma", 0.047);
    }
    const word bk(pbeProperties.lookup("breakupKernel"));
    b200Check(pbe_breakup_type(bk.c_str(), &s.breakup), "breakupKernel::New");
    const dictionary &bd = pbeProperties.optionalSubDict(bk + "Coeffs");
    if (s.breakup == PBE_BR_CT) { // breakupKernels/CoulaloglouTavlarides.C:46-49
        s.br_c[0] = bd.lookupOrDefault<scalar>("c1", 0.00487);
        s.br_c[1] = bd.lookupOrDefault<scalar>("c2", 0.008);
        s.br_c[2] = bd.lookupOrDefault<scalar>("gamma", 0.0);
        s.br_c[3] = bd.lookupOrDefault<scalar>("sigma", 0.047);
    }
    const word dk(pbeProperties.lookup("daughterParticleDistribution"));
    b200Check(pbe_daughter_type(dk.c_str(), &s.daughter), "daughterParticleDistribution::New");
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

namespace PBEMethods {

// ---- MOC ------------------------------------------------------------------------------------------
defineTypeNameAndDebug(MOC, 0);
addToRunTimeSelectionTable(PBEMethod, MOC, dictionary);

static inline std::string fieldName(const char *stem, label i) {
    std::ostringstream os;
    os << stem << i;
    return os.str();
}

MOC::MOC(const dictionary &pbeProperties, const phaseModel &phase)
    : PBEMethod(pbeProperties, phase), runTime(phase.mesh().time()), deltaXi_("deltaXi", 0.0),
      numberOfClasses_(0), usingMULES_(false), source_term_relaxation_(1.0), moc_(nullptr), nHost_(nullptr),
      SHost_(nullptr), continuityError_(0.0), b200Mesh_(nullptr), divScheme_(PBE_DIV_UPWIND) {
    const dictionary &MOCDict_ = pbeProperties.subDict("MOCCoeffs"); // MOC.C:51-60
    numberOfClasses_ = readLabel(MOCDict_.lookup("numberOfClasses"));
    deltaXi_ = dimensionedScalar("deltaXi", readScalar(MOCDict_.lookup("xi1")));
    usingMULES_ = MOCDict_.lookupOrDefault<bool>("usingMULES", false);
    source_term_relaxation_ = MOCDict_.lookupOrDefault<scalar>("sourceRelaxation", 1.0);
    Info << "Creating " << numberOfClasses_ << " class" << (numberOfClasses_ > 1 ? "es" : "") << endl;
    classNumberDensity_ = PtrList<volScalarField>(numberOfClasses_);
    forAll(classNumberDensity_, i) // MOC.C:69-77: fields n0 .. n<N-1>, registered on the mesh
        classNumberDensity_.set(i, new volScalarField(fieldName("n", i), phase.mesh()));
    b200Check(pbe_moc_create(ctx_, numberOfClasses_, deltaXi_.value(), &spec_, source_term_relaxation_, &moc_),
              "MOC::MOC");
    const int64_t bytes = (int64_t)sizeof(double) * numberOfClasses_ * phase.size();
    b200Check(pbe_host_alloc(ctx_, bytes, (void **)&nHost_), "MOC::MOC");
    b200Check(pbe_host_alloc(ctx_, bytes, (void **)&SHost_), "MOC::MOC");
    const fvMesh &mesh = phase.mesh();
    if (mesh.hasFaces() && usingMULES_) {
        // fvSchemes divSchemes "div\\(phi,n.*\\)" (the shim reads it from MOCCoeffs); a scheme without a
        // device path is a FatalError listing the valid ones, like any other run-time selection
        const word scheme = MOCDict_.lookupOrDefault<word>("divScheme", "Gauss upwind");
        b200Check(pbe_div_scheme_type(scheme.c_str(), &divScheme_), "MOC::MOC div(phi,n.*)");
        std::vector<int32_t> own(mesh.owner().begin(), mesh.owner().end()),
            nei(mesh.neighbour().begin(), mesh.neighbour().end()),
            bfc(mesh.boundaryFaceCells_.begin(), mesh.boundaryFaceCells_.end());
        b200Check(pbe_mesh_create(ctx_, mesh.nCells(), mesh.nInternalFaces(), own.data(), nei.data(),
                                  (int64_t)bfc.size(), bfc.data(), mesh.V().cdata(), &b200Mesh_),
                  "MOC::MOC mesh");
    }
}

MOC::~MOC() {
    if (nHost_) pbe_host_free(ctx_, nHost_);
    if (SHost_) pbe_host_free(ctx_, SHost_);
    if (b200Mesh_) pbe_mesh_destroy(b200Mesh_);
    if (moc_) pbe_moc_destroy(moc_);
}

// MOC::correct(), MOC.C:159-184.  Replaced: the breakup-rate cache fill (:162-171),
// coalescence_().update() (:173) and the N calls of classSourceTerm(k) (:174-175, i.e.
// :87-157).  Kept: the transport step (:177-180) and the min(n_k) prints (:181-183).
void MOC::correct() {
    const label nCells = phase_.size();
    forAll(classNumberDensity_, k) // each field is contiguous: it becomes row k of the class-major block
        std::memcpy(nHost_ + (size_t)k * nCells, classNumberDensity_[k].internalField().cdata(),
                    sizeof(double) * nCells);
    updatePhaseFields();
    if (b200Mesh_) {
        correctWithTransport();
        return;
    }
    b200Check(pbe_moc_source_host(moc_, nCells, nCells, nHost_, epsilonData(), rhoHost_.cdata(), nuHost_.cdata(),
                                  runTime.deltaT().value(), SHost_),
              "MOC::correct");
#ifdef PBE_B200_REAL_OPENFOAM
    // wrap row k of SHost_ as the field classSourceTerm(k) used to return, then
    // if (usingMULES_) solveWithMULES(S); else solveWithFVM(S);          (MOC.C:177-180, unchanged)
#error "transport step: use the reference's solveWithMULES / solveWithFVM here"
#else
    solveZeroFlux(SHost_);
#endif
    forAll(classNumberDensity_, k) {
        scalar mn = HUGE_VAL;
        const scalarField &f = classNumberDensity_[k].internalField();
        forAll(f, c) mn = f[c] < mn ? f[c] : mn;
        if (k == 0 || k == numberOfClasses_ - 1) Info << "min(n" << k << ") " << mn << endl; // MOC.C:181-183
    }
}

// MOC::correct() with solveWithMULES (MOC.C:198-272) and `Gauss upwind`: one call; the class fields
// travel to the device once, sources and the transport update run there, the new fields come back.
void MOC::correctWithTransport() {
    const label nCells = phase_.size();
    b200Check(pbe_moc_correct_host(moc_, b200Mesh_, divScheme_, nCells, nHost_, epsilonData(), rhoHost_.cdata(),
                                   nuHost_.cdata(), phase_.phi().cdata(), nullptr /* zeroGradient patches */, 0,
                                   runTime.deltaT().value(), nullptr),
              "MOC::correct");
    forAll(classNumberDensity_, k)
        std::memcpy(classNumberDensity_[k].internalField().data_ptr(), nHost_ + (size_t)k * nCells,
                    sizeof(double) * nCells);
    continuityCheck();
    forAll(classNumberDensity_, k) {
        scalar mn = HUGE_VAL;
        const scalarField &f = classNumberDensity_[k].internalField();
        forAll(f, c) mn = f[c] < mn ? f[c] : mn;
        if (k == 0 || k == numberOfClasses_ - 1) Info << "min(n" << k << ") " << mn << endl; // MOC.C:181-183
    }
}

void MOC::continuityCheck() { // MOC.C:254-279: max |sum_k xi_k n_k / V - alpha|
    const label nCells = phase_.size();
    scalarField volumeSum(nCells, 0.0);
    forAll(classNumberDensity_, k) {
        const scalarField &n = classNumberDensity_[k].internalField();
        const scalar xi = deltaXi_.value() * (k + 1);
        forAll(n, c) volumeSum[c] += xi * n[c];
    }
    continuityError_ = 0.0;
    const scalarField &V = phase_.mesh().V();
    forAll(volumeSum, c) {
        const scalar e = std::fabs(volumeSum[c] / V[c] - phase_[c]);
        continuityError_ = e > continuityError_ ? e : continuityError_;
    }
}

void MOC::solveZeroFlux(const double *S) {
    const label nCells = phase_.size();
    const scalar dt = runTime.deltaT().value();
    scalarField volumeSum(nCells, 0.0);
    forAll(classNumberDensity_, k) { // MULES::explicitSolve(1, n_k, phi = 0, Sp = 0, Su = S_k), MOC.C:259-272
        scalarField &n = classNumberDensity_[k].internalField();
        const scalar xi = deltaXi_.value() * (k + 1);
        forAll(n, c) {
            n[c] = n[c] + dt * S[(size_t)k * nCells + c];
            volumeSum[c] += xi * n[c];
        }
    }
    continuityError_ = 0.0; // MOC.C:254-279: max |sum_k xi_k n_k / V - alpha|
    const scalarField &V = phase_.mesh().V();
    forAll(volumeSum, c) {
        const scalar e = std::fabs(volumeSum[c] / V[c] - phase_[c]);
        continuityError_ = e > continuityError_ ? e : continuityError_;
    }
}

// MOC::d(), MOC.C:282-290 with volumeToDiameter (:43-45): number-mean diameter
const volScalarField MOC::d() const {
    volScalarField result("diameter", phase_.mesh(), 0.0, false);
    const label nCells = phase_.size();
    scalarField num(nCells, 0.0), den(nCells, 0.0);
    forAll(classNumberDensity_, k) {
        const scalar dk = std::pow(6.0 / pi * deltaXi_.value() * (k + 1), 1.0 / 3.0);
        const scalarField &n = classNumberDensity_[k].internalField();
        forAll(n, c) {
            num[c] += dk * n[c];
            den[c] += n[c];
        }
    }
    forAll(num, c) result[c] = num[c] / den[c];
    return result;
}

// ---- QMOM -----------------------------------------------------------------------------------------
defineTypeNameAndDebug(QMOM, 0);
addToRunTimeSelectionTable(PBEMethod, QMOM, dictionary);

QMOM::QMOM(const dictionary &pbeProperties, const phaseModel &phase)
    : PBEMethod(pbeProperties, phase), runTime(phase.mesh().time()),
      quadratureOrder_(readLabel(pbeProperties.subDict("QMOMCoeffs").lookup("quadratureOrder"))), // QMOM.C:63
      moments_(2 * quadratureOrder_), qmom_(nullptr), mHost_(nullptr), SHost_(nullptr),
      d_("diameter", phase.mesh(), 0.0) {
    forAll(moments_, k) moments_.set(k, new volScalarField(fieldName("m", k), phase.mesh())); // QMOM.C:72-79
    b200Check(pbe_qmom_create(ctx_, quadratureOrder_, &spec_, &qmom_), "QMOM::QMOM");
    const int64_t bytes = (int64_t)sizeof(double) * moments_.size() * phase.size();
    b200Check(pbe_host_alloc(ctx_, bytes, (void **)&mHost_), "QMOM::QMOM");
    b200Check(pbe_host_alloc(ctx_, bytes, (void **)&SHost_), "QMOM::QMOM");
}

QMOM::~QMOM() {
    if (mHost_) pbe_host_free(ctx_, mHost_);
    if (SHost_) pbe_host_free(ctx_, SHost_);
    if (qmom_) pbe_qmom_destroy(qmom_);
}

// QMOM::correct(), QMOM.C:85-129.  Replaced: the 2N calls of momentSourceTerm(k) (:88-96, i.e.
// :215-267 with one wheeler_inversion per moment index).  Kept on the OpenFOAM side: the NaN
// scrub (:98-110), the transport step (:112-115), the SMALL clamp (:117-123) and d_ (:125).
void QMOM::correct() {
    const label nCells = phase_.size();
    forAll(moments_, k)
        std::memcpy(mHost_ + (size_t)k * nCells, moments_[k].internalField().cdata(), sizeof(double) * nCells);
    updatePhaseFields();
    b200Check(pbe_qmom_source_host(qmom_, nCells, nCells, mHost_, epsilonData(), rhoHost_.cdata(), nuHost_.cdata(),
                                   SHost_),
              "QMOM::correct");
    const scalar dt = runTime.deltaT().value();
    forAll(moments_, k) {
        scalarField &m = moments_[k].internalField();
        forAll(m, c) {
            scalar S = SHost_[(size_t)k * nCells + c];
            if (std::isnan(S)) S = 0.0; // QMOM.C:101-107
#ifdef PBE_B200_REAL_OPENFOAM
#error "transport step: use the reference's solveWithMULES / solveWithFVM here (QMOM.C:112-115)"
#endif
            m[c] = m[c] + dt * S;       // QMOM.C:193-197 with phi = 0
            m[c] = m[c] > SMALL ? m[c] : SMALL; // QMOM.C:117-123
        }
    }
    forAll(d_.internalField(), c) // QMOM.C:125
        d_[c] = std::pow(6.0 / pi * moments_[1][c] / moments_[0][c], 1.0 / 3.0);
}

} // namespace PBEMethods

// ---- PBEDiameter (PBEDiameter.C:46-75, PBEDiameter.H:95-99) -----------------------------------------
diameterModels::PBEDiameter::PBEDiameter(const dictionary &diameterProperties, const phaseModel &phase)
    : method_(PBEMethod::New(diameterProperties, phase)), d_("d." + phase.name(), phase.mesh(), 0.0) {}

void diameterModels::PBEDiameter::correct() {
    method_->correct();
    const volScalarField d(method_->d());
    forAll(d_.internalField(), c) d_[c] = d[c];
}

} // namespace Foam
