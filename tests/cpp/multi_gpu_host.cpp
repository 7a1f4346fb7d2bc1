// multi_gpu_host.cpp -- the multi-GPU host loop of INTEGRATION.md section 3, as a translation unit that the CPU
// suite compiles and links against libmcrt_b200.so (tests/test_bvh_host.py), so the snippet in the document
// cannot drift from include/ptb.h.  Plain C ABI only: no NCCL, CUDA or torch header is needed on the host side.
//
// One process per GPU.  `carry_id` stands for whatever the engine's launcher uses to move 128 bytes from rank 0 to
// the other ranks (MPI_Bcast, a socket, a file); here: a file path given on the command line.
//
// Run (needs GPUs):  multi_gpu_host <rank> <num_ranks> <id-file>      Without a CUDA device it prints NO_DEVICE.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "ptb.h"
#include "rrb.h"

static void ptbCheck(int rc) {
    if (rc) throw std::runtime_error(ptb_last_error());
}
static void rrbCheck(int rc) {
    if (rc) throw std::runtime_error(rrb_last_error());
}

// --- launcher glue (not part of the library): rank 0 writes the id, the others poll for it -------------------------
static void carry_id(const char* path, int rank, unsigned char id[PTB_NCCL_UNIQUE_ID_BYTES]) {
    if (rank == 0) {
        std::string tmp = std::string(path) + ".tmp";
        FILE* f = fopen(tmp.c_str(), "wb");
        if (!f || fwrite(id, 1, PTB_NCCL_UNIQUE_ID_BYTES, f) != PTB_NCCL_UNIQUE_ID_BYTES) throw std::runtime_error("id file");
        fclose(f);
        rename(tmp.c_str(), path);
        return;
    }
    for (;;) {
        FILE* f = fopen(path, "rb");
        if (f) {
            size_t n = fread(id, 1, PTB_NCCL_UNIQUE_ID_BYTES, f);
            fclose(f);
            if (n == PTB_NCCL_UNIQUE_ID_BYTES) return;
        }
    }
}

// --- what RTScene::commit + the render passes become on N GPUs -----------------------------------------------------
struct HostScene {                 // the RTHostScene arrays the engine already has (kernel_data.h:278-294 layouts)
    ptb_scene_desc desc;
    const void* camera;            // RTPinholeCamera, 176 bytes
    ptb_filter_props filter;
    int width, height, max_depth;
};

// Renders `frames` samples per pixel split over the ranks by image tiles and returns the merged image on rank 0.
// `attach` is the engine's own RTScene code that creates the Radeon Rays shapes on `api` and commits.
std::vector<float> render_on_rank(int rank, int num_ranks, const char* id_path, const HostScene& s, int frames,
                                  void (*attach)(rrb_api*)) {
    rrb_api* api = nullptr;
    rrbCheck(rrb_create(rank, nullptr, &api));                    // device `rank`, library-owned stream
    attach(api);                                                  // CreateMesh / AttachShape / Commit: replicated
    ptb_context* ctx = nullptr;
    ptbCheck(ptb_create(api, &ctx));
    ptbCheck(ptb_upload_scene(ctx, &s.desc));                     // replicated scene
    ptbCheck(ptb_resize(ctx, s.width, s.height));
    ptbCheck(ptb_set_partition(ctx, rank, num_ranks));            // 64x64 tiles, round-robin; full-frame buffers

    unsigned char id[PTB_NCCL_UNIQUE_ID_BYTES];
    if (rank == 0) ptbCheck(ptb_nccl_unique_id(id));
    carry_id(id_path, rank, id);
    void* comm = nullptr;
    ptbCheck(ptb_nccl_comm_create(ctx, id, rank, num_ranks, &comm));

    const size_t P = (size_t)s.width * s.height;
    std::vector<float> image;
    for (int f = 0; f < frames; ++f) {
        ptbCheck(ptb_set_camera(ctx, s.camera));                  // per-frame TAA offset goes in here
        ptbCheck(ptb_pt_render_sample(ctx, f, s.max_depth));      // this rank's tiles only
        ptbCheck(ptb_accumulate(ctx, f, &s.filter));              // rank-local sum(w L) || sum(w)
    }
    // the ONE collective of the path: ncclReduce of 20 B/pixel to rank 0 on the context's stream + the division
    ptbCheck(ptb_merge_accumulation(ctx, comm, /*root=*/0));
    if (rank == 0) {
        image.resize(4 * P);
        ptbCheck(ptb_read_buffer(ctx, "NextFrameImageCL", image.data(), image.size() * sizeof(float)));
    } else {
        ptbCheck(ptb_synchronize(ctx));
    }
    ptbCheck(ptb_nccl_comm_destroy(comm));
    ptbCheck(ptb_destroy(ctx));
    rrbCheck(rrb_destroy(api));
    return image;
}

int main(int argc, char** argv) {
    // The CPU suite only needs this file to compile and link; run without arguments it checks that the library
    // refuses to work without a device instead of falling back.
    rrb_api* api = nullptr;
    int rc = rrb_create(0, nullptr, &api);
    if (rc) {
        printf("NO_DEVICE rrb_create -> %d: %s\n", rc, rrb_last_error());
        return 0;
    }
    rrb_destroy(api);
    printf("OK device present; render_on_rank is exercised by tests/nccl_merge_worker.py through the same entry points (%d args)\n", argc);
    (void)argv;
    return 0;
}
