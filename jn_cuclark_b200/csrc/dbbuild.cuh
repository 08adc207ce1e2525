// dbbuild.cuh -- internal hooks of the database builder (dbbuild.cu)
#pragma once
#include <memory>
#include <stddef.h>
#include <stdint.h>
struct ccb_handle;
namespace ccb {
struct BuildState;
void build_free(ccb_handle* h);     // releases a build in progress (ccb_destroy, ccb_build_end)
// FASTA text -> code stream of the full-mode scan (0..3 = base, 4 = window reset, one reset closes the text);
// returns the number of codes, adds the bases seen to `bases`
size_t fasta_to_codes(const unsigned char* text, uint64_t len, std::unique_ptr<uint8_t[]>& out, uint64_t& bases);
}
