// See wenoB200Registry.H.
#include "wenoB200Registry.H"
#include "IOdictionary.H"
#include "Pstream.H"
#include "OSspecific.H"

#include <map>
#include <utility>
#include <cstdlib>
#include <cstring>

namespace Foam
{
namespace wenoB200
{

namespace
{
    struct Registry
    {
        std::map<std::pair<const fvMesh*, label>, PlanEntry> plans;
        ~Registry()
        {
            for (auto& kv : plans) kv.second.release();
        }
    };
    Registry& theRegistry()
    {
        static Registry r;
        return r;
    }
}


double* PlanEntry::reserve(double*& buf, size_t& cap, const size_t n)
{
    if (cap < n || !buf)
    {
        if (buf) wenoext_host_free(buf);
        buf = nullptr;
        cap = 0;
        void* p = nullptr;
        if (wenoext_host_alloc((n ? n : 1)*sizeof(double), &p) != WENOEXT_OK)
        {
            FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
        }
        buf = static_cast<double*>(p);
        cap = n;
    }
    return buf;
}


void PlanEntry::release()
{
    if (plan) wenoext_plan_destroy(plan);
    if (flux) wenoext_host_free(flux);
    if (pnf) wenoext_host_free(pnf);
    if (corr) wenoext_host_free(corr);
    plan = nullptr;
    flux = pnf = corr = nullptr;
    fluxCap = pnfCap = corrCap = 0;
}


void clearRegistry()
{
    for (auto& kv : theRegistry().plans) kv.second.release();
    theRegistry().plans.clear();
}


int deviceOfThisRank()
{
    // launchers export the local rank; fall back to rank modulo visible devices
    for (const char* name : {"OMPI_COMM_WORLD_LOCAL_RANK", "MV2_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID"})
    {
        if (const char* v = std::getenv(name)) return std::atoi(v);
    }
    const int n = wenoext_device_count();
    if (n < 1)
    {
        FatalErrorInFunction << "wenoext_b200: no CUDA device (the library has no CPU fallback)"
            << exit(FatalError);
    }
    return Pstream::parRun() ? Pstream::myProcNo() % n : 0;
}


PlanEntry& registry(const fvMesh& mesh, const WENOBase& base, const label polOrder)
{
    auto& plans = theRegistry().plans;
    const auto key = std::make_pair(&mesh, polOrder);
    auto it = plans.find(key);
    if (it != plans.end()) return it->second;

    // system/WENODict: the reference's keys with the reference's defaults (WENOCoeff.C:41-48, :87-89)
    // opened exactly as the reference does (WENOCoeff.C:70-80)
    IOdictionary WENODict
    (
        IOobject
        (
            "WENODict",
            mesh.time().caseSystem(),
            mesh,
            IOobject::READ_IF_PRESENT,
            IOobject::NO_WRITE
        )
    );
    const scalar p   = WENODict.lookupOrDefault<scalar>("p", 4.0);
    const scalar dm  = WENODict.lookupOrDefault<scalar>("dm", 1000.0);
    const scalar eps = WENODict.lookupOrDefault<scalar>("epsilon", 1e-40);
    const bool useCache = WENODict.lookupOrDefault<bool>("useCache", false);

    const int device = deviceOfThisRank();
    const fileName cacheDir = mesh.time().path()/"constant"/("WENOBase" + Foam::name(polOrder));
    wenoext_plan* plan =
        (useCache && isDir(cacheDir)) ? planFromCache(mesh, polOrder, device)
                                      : planFromBase(mesh, base, polOrder, device);

    if (wenoext_set_params(plan, p, dm, eps) != WENOEXT_OK)
    {
        FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
    }

    if (Pstream::parRun())
    {
        // rank 0 creates the id, everybody gets it over Pstream, then all ranks join
        // (collectData and swapData of the reference become the library's two exchanges)
        List<char> id(128, '\0');
        if (Pstream::master() && wenoext_nccl_unique_id(id.data()) != WENOEXT_OK)
        {
            FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
        }
        Pstream::scatter(id);
        if (wenoext_attach_nccl(plan, id.cdata(), Pstream::myProcNo(), Pstream::nProcs()) != WENOEXT_OK)
        {
            FatalErrorInFunction << "wenoext_b200: " << wenoext_last_error() << exit(FatalError);
        }
    }

    PlanEntry& e = plans[key];
    e.plan = plan;
    return e;
}

}
}
