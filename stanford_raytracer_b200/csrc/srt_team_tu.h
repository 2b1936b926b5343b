// srt_team_tu.h -- entry points of the team-kernel translation unit (srt_team_tu.cu)
#pragma once
#include <cuda_runtime.h>
struct SrtModel;
struct SrtRunParams;
struct SrtBatch;
int srt_team_block_threads(long long nrays, int num_sms);
cudaError_t srt_team_occupancy(const SrtModel& M, int block, int* blocks_per_sm);
cudaError_t srt_team_attributes(const SrtModel& M, int block, cudaFuncAttributes* fa);
// the model of this unit's kernels lives in its own __constant__ copy: upload, then launch on `stream`
cudaError_t srt_team_upload_model(const SrtModel& M, cudaStream_t stream);
cudaError_t srt_team_launch(const SrtModel& M, const SrtRunParams& P, const SrtBatch& B, int block, int grid, cudaStream_t stream);
