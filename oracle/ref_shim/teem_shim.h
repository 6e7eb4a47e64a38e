/*
 * teem_shim.h -- TEST INFRASTRUCTURE.  The four teem calls photon's loadNRRD makes
 * (trace_rays_through_density_gradients.h:1676-1814: nrrdNew, nrrdLoad, biffGetDone, nrrdNuke;
 * fields axis[i].size, axis[i].spacing, spaceOrigin[i], data) on top of oracle/nrrd_read.h.
 * teem itself ships only as an un-loadable binary in the reference.
 */
#ifndef PHOTON_REF_TEEM_SHIM_H_
#define PHOTON_REF_TEEM_SHIM_H_
#include "../nrrd_read.h"

#define NRRD 0
struct NrrdAxisInfo { size_t size; double spacing; };
struct Nrrd {
    NrrdAxisInfo axis[16];
    double spaceOrigin[8];
    void *data;
    onrrd::Volume *vol;
};
namespace shim { extern thread_local std::string t_biff; }
inline Nrrd *nrrdNew() { Nrrd *n = new Nrrd(); memset(n, 0, sizeof *n); return n; }
inline int nrrdLoad(Nrrd *n, const char *path, void *) {
    n->vol = new onrrd::Volume();
    if (!onrrd::load(path, *n->vol)) { shim::t_biff = n->vol->error; return 1; }
    n->axis[0].size = (size_t)n->vol->nx; n->axis[1].size = (size_t)n->vol->ny; n->axis[2].size = (size_t)n->vol->nz;
    for (int i = 0; i < 3; i++) { n->axis[i].spacing = n->vol->spacing[i]; n->spaceOrigin[i] = n->vol->origin[i]; }
    n->data = n->vol->rho.data();
    return 0;
}
inline char *biffGetDone(int) { return const_cast<char *>(shim::t_biff.c_str()); }
inline Nrrd *nrrdNuke(Nrrd *n) { if (n) { delete n->vol; delete n; } return nullptr; }
#endif
