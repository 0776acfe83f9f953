// correction() of the drop-in scheme classes (WENOUpwindFitB200.H): what replaces the body of
// WENOUpwindFit<Type>::correction (libWENOEXT/WENOUpwindFit/WENOUpwindFit.C:212-290) and of
// WENOUpwindFit01<Type>::correction (libWENOEXT/WENOUpwindFit01/WENOUpwindFit01.C:215-292) by one call
// into libwenoext_b200.so. See INTEGRATION.md.
#include <cstring>

template<class Type, bool Scheme01>
Foam::tmp<Foam::GeometricField<Type, Foam::fvsPatchField, Foam::surfaceMesh> >
Foam::WENOUpwindFitB200Base<Type, Scheme01>::correction
(
    const GeometricField<Type, fvPatchField, volMesh>& vf
) const
{
    const fvMesh& mesh = this->mesh();

    // creates the plan on first use (WENODict factors, NCCL attach in a parallel run)
    wenoB200::PlanEntry& entry = wenoB200::registry(mesh, WENOBase_, polOrder_);

    // the result field exactly as the reference builds it (WENOUpwindFit.C:230-246): named "tvfP",
    // not registered, dimensions of vf, zero everywhere (so non-coupled patches stay zero)
    tmp<GeometricField<Type, fvsPatchField, surfaceMesh> > tsfCorrP
    (
        new GeometricField<Type, fvsPatchField, surfaceMesh>
        (
            IOobject
            (
                "tvfP",
                mesh.time().timeName(),
                mesh,
                IOobject::NO_READ,
                IOobject::NO_WRITE,
                false
            ),
            mesh,
            dimensioned<Type>(vf.name(), vf.dimensions(), pTraits<Type>::zero)
        )
    );
    GeometricField<Type, fvsPatchField, surfaceMesh>& tsfP =
    #ifdef FOAM_NEW_TMP_RULES
        tsfCorrP.ref();
    #else
        tsfCorrP();
    #endif

    // AoS buffers in the ABI's order: internal faces first, then the patches in mesh.boundary() order.
    // Type[n] is n*nComp contiguous doubles (-DWM_DP), so fields are copied as raw memory.
    const label nInt = mesh.nInternalFaces();
    const label nFaces = mesh.nFaces();
    const int nComp = pTraits<Type>::nComponents;
    const bool needNbr = limFac_ && !Scheme01;       // only WENOUpwindFit's limiter reads it (WENOUpwindFit.C:130)

    double* flux = wenoB200::PlanEntry::reserve(entry.flux, entry.fluxCap, size_t(nFaces));
    double* corr = wenoB200::PlanEntry::reserve(entry.corr, entry.corrCap, size_t(nFaces)*nComp);
    double* pnf = nullptr;

    const Field<scalar>& fluxI =
    #ifdef FOAM_NEW_GEOMFIELD_RULES
        faceFlux_.primitiveField();
    #else
        faceFlux_.internalField();
    #endif
    std::memcpy(flux, fluxI.cdata(), size_t(nInt)*sizeof(double));

    const fvPatchList& patches = mesh.boundary();
    forAll(patches, patchI)
    {
        const fvPatch& fp = patches[patchI];
        const fvsPatchField<scalar>& pFlux = faceFlux_.boundaryField()[patchI];
        if (pFlux.size())
        {
            std::memcpy(flux + fp.start(), pFlux.cdata(), size_t(pFlux.size())*sizeof(double));
        }
        if (needNbr && fp.coupled() && fp.size())
        {
            if (!pnf)
            {
                // values on faces of non-coupled patches are never read
                pnf = wenoB200::PlanEntry::reserve(entry.pnf, entry.pnfCap, size_t(nFaces - nInt)*nComp);
            }
            const tmp<Field<Type> > tnbr(vf.boundaryField()[patchI].patchNeighbourField());
            const Field<Type>& nbr = tnbr();
            std::memcpy
            (
                pnf + size_t(fp.start() - nInt)*nComp,
                nbr.cdata(),
                size_t(nbr.size())*sizeof(Type)
            );
        }
    }

    const Field<Type>& vfI =
    #ifdef FOAM_NEW_GEOMFIELD_RULES
        vf.primitiveField();
    #else
        vf.internalField();
    #endif

    const int rc = wenoext_correction
    (
        entry.plan,
        Scheme01 ? WENOEXT_SCHEME_UPWINDFIT01 : WENOEXT_SCHEME_UPWINDFIT,
        int(limFac_),
        nComp,
        reinterpret_cast<const double*>(vfI.cdata()),
        flux,
        pnf,
        corr,
        WENOEXT_MEM_HOST,
        nullptr
    );
    if (rc != WENOEXT_OK)
    {
        FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
    }

    Field<Type>& out =
    #ifdef FOAM_NEW_GEOMFIELD_RULES
        tsfP.primitiveFieldRef();
    #else
        tsfP.internalField();
    #endif
    std::memcpy(static_cast<void*>(out.data()), corr, size_t(nInt)*sizeof(Type));
    forAll(patches, patchI)
    {
        fvsPatchField<Type>& pf =
        #ifdef FOAM_NEW_GEOMFIELD_RULES
            tsfP.boundaryFieldRef()[patchI];
        #else
            tsfP.boundaryField()[patchI];
        #endif
        if (pf.size() && patches[patchI].coupled())
        {
            std::memcpy(static_cast<void*>(pf.data()), corr + size_t(patches[patchI].start())*nComp, size_t(pf.size())*sizeof(Type));
        }
    }
    return tsfCorrP;
}
