// dbbuild.cuh -- internal hooks of the database builder (dbbuild.cu)
#pragma once
struct ccb_handle;
namespace ccb {
struct BuildState;
void build_free(ccb_handle* h);     // releases a build in progress (ccb_destroy, ccb_build_end)
}
