// Run-time selection entries of both schemes for the five field types, as
// libWENOEXT/WENOUpwindFit/makeWENOUpwindFit.C:37 and makeWENOUpwindFit01.C do for the reference
// classes. Build this file INSTEAD of the reference's two make*.C files,
// so that "WENOUpwindFit" / "WENOUpwindFit01" in fvSchemes select the B200 classes.
#include "WENOUpwindFitB200.H"
#include "fvMesh.H"

namespace Foam
{
    makeSurfaceInterpolationScheme(WENOUpwindFit)
    makeSurfaceInterpolationScheme(WENOUpwindFit01)
}
