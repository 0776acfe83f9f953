// correction() of the drop-in scheme classes (WENOUpwindFitB200.H): what replaces the body of
// WENOUpwindFit<Type>::correction (libWENOEXT/WENOUpwindFit/WENOUpwindFit.C:212-290) and of
// WENOUpwindFit01<Type>::correction (libWENOEXT/WENOUpwindFit01/WENOUpwindFit01.C:215-292) by one call
// into libwenoext_b200.so. SOURCE ONLY (needs OpenFOAM). See INTEGRATION.md.
template<class Type, bool Scheme01>
Foam::tmp<Foam::GeometricField<Type, Foam::fvsPatchField, Foam::surfaceMesh>>
Foam::WENOUpwindFitB200Base<Type, Scheme01>::correction
(
    const GeometricField<Type, fvPatchField, volMesh>& vf
) const
{
    const fvMesh& mesh = this->mesh();
    wenoext_plan* plan = wenoB200::registry(mesh, WENOBase_, polOrder_);   // creates on first use, calls
                                                                           // wenoext_set_params with WENODict's
                                                                           // p / dm / epsilon and attaches NCCL
    tmp<GeometricField<Type, fvsPatchField, surfaceMesh>> tsfP(/* "tvfP", vf.dimensions(), zero: as :230-246 */);
    auto& sf = tsfP.ref();

    // AoS host buffers in the ABI's order: internal faces first, then the patches in mesh.boundary() order
    const label nInt = mesh.nInternalFaces(), nFaces = mesh.nFaces();
    constexpr int nComp = pTraits<Type>::nComponents;
    static thread_local std::vector<double> flux, pnf, corr;             // pinned in a real shim (cudaHostAlloc)
    flux.resize(nFaces); pnf.assign(size_t(nFaces - nInt)*nComp, 0.0); corr.resize(size_t(nFaces)*nComp);
    std::copy(faceFlux_.primitiveField().begin(), faceFlux_.primitiveField().end(), flux.begin());
    forAll(mesh.boundary(), p)
    {
        const fvPatch& fp = mesh.boundary()[p];
        const auto& pf = faceFlux_.boundaryField()[p];
        for (label i = 0; i < pf.size(); ++i) flux[fp.start() + i] = pf[i];
        if (limFac_ && vf.boundaryField()[p].coupled())                  // WENOUpwindFit.C:118-165
        {
            const Field<Type> nb(vf.boundaryField()[p].patchNeighbourField());
            std::memcpy(&pnf[size_t(fp.start() - nInt)*nComp], nb.cdata(), nb.size()*sizeof(Type));
        }
    }
    const int rc = wenoext_correction
    (
        plan, Scheme01 ? WENOEXT_SCHEME_UPWINDFIT01 : WENOEXT_SCHEME_UPWINDFIT, int(limFac_), nComp,
        reinterpret_cast<const double*>(vf.primitiveField().cdata()),    // Type[nCells] is AoS fp64 (-DWM_DP)
        flux.data(), limFac_ ? pnf.data() : nullptr, corr.data(), WENOEXT_MEM_HOST, nullptr
    );
    if (rc != WENOEXT_OK) FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);

    std::memcpy(sf.primitiveFieldRef().data(), corr.data(), size_t(nInt)*sizeof(Type));
    forAll(mesh.boundary(), p)
    {
        auto& pf = sf.boundaryFieldRef()[p];
        if (pf.size()) std::memcpy(pf.data(), &corr[size_t(mesh.boundary()[p].start())*nComp], pf.size()*sizeof(Type));
    }
    return tsfP;
}
