// shimMain.C -- drives the adapter the way twoPhaseEulerFoam drives libPBE.so: construct the
// diameterModel from a PBEDiameterCoeffs dictionary (run-time selection by `method`), then per time
// step  diameterModel.correct()  (PBEDiameter.H:95-99; applications/PBEFoam/PBEFoam.C:56-67).
// Fields come from a raw FP64 file so that tests/test_adapter_shim.py can compare the result with
// the oracle.  Shim build only (adapter/foam_shim): test infrastructure, not the product.
//
//   shimMain <dict file> <input.bin> <output.bin> <steps> [mesh.bin]
//   dict  : the PBEDiameterCoeffs dictionary; the driver wraps it as the phase's `diameterModel PBEDiameter;
//           PBEDiameterCoeffs { ... }` and goes through diameterModel::New, as the solver's phaseModel does
//   mesh  : [nInternal, nBoundary] (FP64), owner[nInternal], neighbour[nInternal], faceCells[nBoundary]
//           (int32), phi[nInternal + nBoundary] (FP64): face addressing and the phase's face flux; optionally
//           followed by Sf[faces][3], C[cells][3], weights[nInternal] (FP64) for the limited schemes
//   input : [K, nCells, dt, cellVolume] then state[K][nCells], epsilon, rho, nu, alpha [nCells] each
//           (K = numberOfClasses for methodOfClasses, 2 quadratureOrder for QMOM)
//   output: state[K][nCells], d[nCells], continuity error (MOC; 0 for QMOM)
#include <cstdio>
#include <fstream>
#include <iterator>

#include "PBEMethodB200.H"

using namespace Foam;

int main(int argc, char **argv) {
    if (argc < 5) {
        std::fprintf(stderr, "usage: shimMain <dict> <input.bin> <output.bin> <steps>\n");
        return 2;
    }
    try {
        std::ifstream df(argv[1]);
        const std::string text((std::istreambuf_iterator<char>(df)), std::istreambuf_iterator<char>());
        const dictionary dict = dictionary::fromString(text, "PBEDiameterCoeffs");
        std::FILE *f = std::fopen(argv[2], "rb");
        double hdr[4];
        if (!f || std::fread(hdr, 8, 4, f) != 4) return 3;
        const label K = (label)hdr[0], nCells = (label)hdr[1];
        std::vector<double> state((size_t)K * nCells), eps(nCells), rho(nCells), nu(nCells), alpha(nCells);
        bool ok = std::fread(state.data(), 8, state.size(), f) == state.size();
        for (std::vector<double> *v : {&eps, &rho, &nu, &alpha}) ok = ok && std::fread(v->data(), 8, nCells, f) == (size_t)nCells;
        std::fclose(f);
        if (!ok) return 3;

        Time runTime(hdr[2]);
        fvMesh mesh(runTime, nCells, hdr[3]);
        if (argc > 5) {
            std::FILE *mf = std::fopen(argv[5], "rb");
            double mh[2];
            if (!mf || std::fread(mh, 8, 2, mf) != 2) return 3;
            const size_t ni = (size_t)mh[0], nb = (size_t)mh[1];
            std::vector<int32_t> own(ni), nei(ni), bfc(nb);
            bool mok = std::fread(own.data(), 4, ni, mf) == ni && std::fread(nei.data(), 4, ni, mf) == ni &&
                       std::fread(bfc.data(), 4, nb, mf) == nb;
            mesh.setFaces(std::vector<label>(own.begin(), own.end()), std::vector<label>(nei.begin(), nei.end()),
                          std::vector<label>(bfc.begin(), bfc.end()));
            mok = mok && std::fread(mesh.phi_.data_ptr(), 8, ni + nb, mf) == ni + nb;
            std::vector<double> Sf(3 * (ni + nb)), Cc(3 * (size_t)nCells), wts(ni);
            if (mok && std::fread(Sf.data(), 8, Sf.size(), mf) == Sf.size() && std::fread(Cc.data(), 8, Cc.size(), mf) == Cc.size() &&
                std::fread(wts.data(), 8, ni, mf) == ni) {
                mesh.Sf_ = Sf;
                mesh.C_ = Cc;
                mesh.weights_ = wts;
            }
            std::fclose(mf);
            if (!mok) return 3;
        }
        phaseModel phase(mesh, "air", 0.0, 0.0, 0.0);
        volScalarField epsilon("epsilon." + phase.name(), mesh);
        for (label c = 0; c < nCells; ++c) {
            phase[c] = alpha[c];
            phase.rhoField()[c] = rho[c];
            phase.nuField()[c] = nu[c];
            epsilon[c] = eps[c];
        }
        // twoPhaseEulerFoam: phaseModel's constructor calls diameterModel::New(phaseDict_, *this)
        const dictionary phaseDict =
            dictionary::fromString("diameterModel PBEDiameter; PBEDiameterCoeffs { " + text + " }", "phaseProperties.air");
        autoPtr<diameterModel> dPtr = diameterModel::New(phaseDict, phase);
        diameterModels::PBEDiameter &diameter = dynamic_cast<diameterModels::PBEDiameter &>(dPtr());
        PBEMethods::MOC *moc = dynamic_cast<PBEMethods::MOC *>(&diameter.method());
        PBEMethods::QMOM *qmom = dynamic_cast<PBEMethods::QMOM *>(&diameter.method());
        PBEMethods::MOM *mom = dynamic_cast<PBEMethods::MOM *>(&diameter.method());
        auto field = [&](label k) -> volScalarField * {
            return moc ? &moc->classNumberDensity(k) : qmom ? &qmom->moment(k) : mom ? &mom->moment(k) : nullptr;
        };
        for (label k = 0; k < K; ++k)
            if (volScalarField *fld = field(k))
                for (label c = 0; c < nCells; ++c) (*fld)[c] = state[(size_t)k * nCells + c];
        const int steps = std::atoi(argv[4]);
        runTime.setWriteInterval(steps > 0 ? steps : 1); // the last step is a write step: fields are synchronised
        for (int s = 0; s < steps; ++s) {
            runTime++;
            diameter.correct();
        }
        for (label k = 0; k < K; ++k)
            if (volScalarField *fld = field(k))
                for (label c = 0; c < nCells; ++c) state[(size_t)k * nCells + c] = (*fld)[c];
        f = std::fopen(argv[3], "wb");
        std::fwrite(state.data(), 8, state.size(), f);
        std::fwrite(diameter.d()().internalField().cdata(), 8, nCells, f);
        const double ce = moc ? moc->continuityError() : 0.0;
        std::fwrite(&ce, 8, 1, f);
        std::fclose(f);
    } catch (const FatalErrorException &e) {
        std::cout << e.what() << "\n\nFOAM exiting\n" << std::endl;
        return 1;
    }
    return 0;
}
