// ref_graphio_driver.cu -- thin command-line driver around the REFERENCE's own
// host-side graph code, compiled from the sources where they lie under
// /root/reference (see oracle/Makefile; output goes to oracle/_ref/ only).
//
// TEST INFRASTRUCTURE ONLY: used to validate oracle/oracle.c's restatement of
// the reader / COO->CSR canonicalisation / source selection and to generate
// the fixtures under tests/golden/ (tests/golden/make_golden.py).
//
// Nothing here is reference code: it only #includes the reference headers
//   gunrock/coo.cuh, gunrock/csr.cuh, gunrock/graphio/{market,rmat}.cuh
// and calls their public functions.
//
// usage:
//   ref_graphio market <file.mtx> <undirected 0|1> <out.bin>
//   ref_graphio rmat <scale> <edgefactor> <srand-seed> <undirected 0|1> <out.bin>
//   ref_graphio randomnode <n> [count]
// out.bin: int64 n, int64 m, int32 has_values, int32 row_offsets[n+1],
//          int32 col[m], int32 values[m] (if has_values)
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <gunrock/coo.cuh>
#include <gunrock/csr.cuh>
#include <gunrock/graphio/market.cuh>
#include <gunrock/graphio/rmat.cuh>

typedef gunrock::Csr<int, int, int> RefCsr;

static int dump(const RefCsr &g, const char *path, bool values)
{
    FILE *f = fopen(path, "wb");
    if (!f) { perror("open output"); return 1; }
    int64_t n = g.nodes, m = g.edges;
    int32_t hv = (values && g.edge_values) ? 1 : 0;
    fwrite(&n, 8, 1, f); fwrite(&m, 8, 1, f); fwrite(&hv, 4, 1, f);
    fwrite(g.row_offsets, 4, (size_t)n + 1, f);
    fwrite(g.column_indices, 4, (size_t)m, f);
    if (hv) fwrite(g.edge_values, 4, (size_t)m, f);
    fclose(f);
    return 0;
}

int main(int argc, char **argv)
{
    if (argc >= 5 && !strcmp(argv[1], "market")) {
        RefCsr g(false);
        if (gunrock::graphio::BuildMarketGraph<true>(argv[2], g, atoi(argv[3]) != 0, false) != 0)
            return 2;
        return dump(g, argv[4], true);
    }
    if (argc >= 7 && !strcmp(argv[1], "rmat")) {
        int scale = atoi(argv[2]), ef = atoi(argv[3]);
        srand((unsigned)atoi(argv[4]));
        RefCsr g(false);
        if (gunrock::graphio::BuildRmatGraph<false>(1 << scale, ef << scale, g, atoi(argv[5]) != 0) != 0)
            return 2;
        return dump(g, argv[6], false);
    }
    if (argc >= 3 && !strcmp(argv[1], "randomnode")) {
        int n = atoi(argv[2]), count = argc >= 4 ? atoi(argv[3]) : 1;
        for (int i = 0; i < count; ++i) printf("%d\n", gunrock::graphio::RandomNode(n));
        return 0;
    }
    fprintf(stderr, "usage: see header of ref_graphio_driver.cu\n");
    return 64;
}
