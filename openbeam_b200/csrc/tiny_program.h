// tiny_program.h — flat layout of a shared-topology "program" (host builder + device reader).
#pragma once
#include "obgpu_internal.h"

namespace obgpu {

#ifdef __CUDACC__
#define OB_HD __host__ __device__
#else
#define OB_HD
#endif

// ints:  [0..7] header N,E,n,nf,nb,NL,EL,nfp | enode[2E] | etype[E] | udof[6E] | edst[21E] | elptr[E+1] |
//        elslot[EL] | nldst[NL] | freek[nf] | bndk[nb] | elkind[EL]
// dbls:  xy[2N] | eai[3E] | nlval[NL] | elq[EL] | eldir[3EL] | elq2[EL]
// (elkind / elq2: OBGPU_LOAD_* and second parameter of every element load; only the run-time specialised kernel reads
// them — the precompiled interpreter handles uniform loads only)
struct TinyLayout {
  int o_enode, o_etype, o_udof, o_edst, o_elptr, o_elslot, o_nldst, o_freek, o_bndk, o_elkind, n_ints;
  int o_xy, o_eai, o_nlval, o_elq, o_eldir, o_elq2, n_dbls;
  OB_HD TinyLayout(int N, int E, int n, int nf, int nb, int NL, int EL) {
    int p = 8;
    o_enode = p; p += 2 * E;
    o_etype = p; p += E;
    o_udof = p; p += 6 * E;
    o_edst = p; p += ELEM_K * E;
    o_elptr = p; p += E + 1;
    o_elslot = p; p += EL;
    o_nldst = p; p += NL;
    o_freek = p; p += nf;
    o_bndk = p; p += nb;
    o_elkind = p; p += EL;
    n_ints = (p + 1) & ~1;  // keep the double section 8-byte aligned when both live in one smem block
    int q = 0;
    o_xy = q; q += 2 * N;
    o_eai = q; q += 3 * E;
    o_nlval = q; q += NL;
    o_elq = q; q += EL;
    o_eldir = q; q += 3 * EL;
    o_elq2 = q; q += EL;
    n_dbls = q;
    (void)n;
  }
};

}  // namespace obgpu
