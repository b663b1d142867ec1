// foamShim.C -- the one definition the header-only API shim needs: diameterModel's TypeName (in a real
// build it lives in OpenFOAM's twoPhaseSystem library).  Shim build only: test infrastructure.
#include "foamShim.H"

namespace Foam {
defineTypeNameAndDebug(diameterModel, 0);
}
