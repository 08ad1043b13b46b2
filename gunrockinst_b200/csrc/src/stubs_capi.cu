// stubs_capi.cu -- gunrock.h entry points outside the accelerated path (SURVEY.md section 8f,
// rank 1): exported so that every program under the reference's shared_lib_tests/ links against
// this library; they report that the primitive is not part of this build and return without
// touching their outputs (reference implementations: gunrock/app/bc/bc_app.cu,
// gunrock/app/topk/topk_app.cu).
#include <cstdio>
#include "../../../include/gunrock/gunrock.h"

extern "C" void gunrock_bc_func(struct GunrockGraph *graph_out, const struct GunrockGraph *graph_in,
                                struct GunrockConfig configs, struct GunrockDataType data_type)
{
    (void)graph_out; (void)graph_in; (void)configs; (void)data_type;
    fprintf(stderr, "libgunrock (B200 build): gunrock_bc_func is outside the accelerated advance/filter path "
                    "and is not built\n");
}

extern "C" void gunrock_topk_func(struct GunrockGraph *graph_out, void *node_ids, void *in_degrees, void *out_degrees,
                                  const struct GunrockGraph *graph_in, struct GunrockConfig configs,
                                  struct GunrockDataType data_type)
{
    (void)graph_out; (void)node_ids; (void)in_degrees; (void)out_degrees; (void)graph_in; (void)configs; (void)data_type;
    fprintf(stderr, "libgunrock (B200 build): gunrock_topk_func is outside the accelerated advance/filter path "
                    "and is not built\n");
}
