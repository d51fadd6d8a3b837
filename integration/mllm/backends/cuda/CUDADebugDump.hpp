// mllm/backends/cuda/CUDADebugDump.hpp -- debugging aid, compiled in only by `make -C integration DEBUG_DUMP=1`: with
// MLLM_DUMP_OPS=<dir> every op's fp32 outputs are written in logical (b, s, h, d) order as <dir>/<n>_<tensor name>.bin
// ([b, h, s, d] int32 header + data), by the CUDA backend's runOp and -- through a one-line sed in the Makefile -- by the CPU
// backend's, so the two backends can be compared op by op on the same graph (profiles/compare_op_dumps.py).
#pragma once
#include "Tensor.hpp"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

namespace mllm {
inline void mllm_debug_dump(std::vector<std::shared_ptr<Tensor>> &outs) {
    static const char *dir = getenv("MLLM_DUMP_OPS");
    if (!dir) return;
    static const int max_rows = getenv("MLLM_DUMP_MAXROWS") ? atoi(getenv("MLLM_DUMP_MAXROWS")) : 1 << 30; // (b, s, h) rows kept per tensor
    static std::map<std::string, int> seen;
    for (auto &t : outs) {
        if (!t || t->dtype() != MLLM_TYPE_F32 || t->shape().size() != 4 || t->count() == 0 || t->rawHostPtr() == nullptr) continue;
        std::string nm = t->name();
        for (auto &c : nm) if (c == '/' || c == ' ') c = '_';
        const int k = seen[nm]++;
        char path[1024];
        snprintf(path, sizeof path, "%s/%s@%d.bin", dir, nm.c_str(), k);
        FILE *f = fopen(path, "wb");
        if (!f) continue;
        int hdr[4] = {t->batch(), t->head(), t->sequence(), t->dimension()};
        // with a row limit: the LAST positions of the sequence (the logits row is among them)
        const int keep_s = std::max(1, std::min(hdr[2], max_rows / std::max(1, hdr[0] * hdr[1]))), s0 = hdr[2] - keep_s;
        hdr[2] = keep_s;
        fwrite(hdr, 4, 4, f);
        std::vector<float> row(hdr[3]);
        for (int b = 0; b < hdr[0]; b++)
            for (int s = s0; s < s0 + keep_s; s++)
                for (int h = 0; h < hdr[1]; h++) {
                    for (int d = 0; d < hdr[3]; d++) row[d] = t->dataAt<float>(b, h, s, d);
                    fwrite(row.data(), 4, row.size(), f);
                }
        fclose(f);
    }
}
} // namespace mllm
