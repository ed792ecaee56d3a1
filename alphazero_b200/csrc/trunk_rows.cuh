// k_net_trunk_rows (trunk_rows_kernels.cu): the convolution trunk for 8 x 8 boards in batches of 1024 positions and
// more (tile = one board row of 14 boards, one TMEM accumulator per output tile); launched by az_net_trunk.
#pragma once
#include <stdint.h>

#define AZ_TRUNK_ROWS_BOARDS_PER_GROUP 14
#define AZ_TRUNK_ROWS_BLOB_OFFSET 56832   // bytes of k_net_trunk's weight blob, which the kernel's own section follows
#define AZ_TRUNK_ROWS_BLOB_BYTES 55296    // conv2..conv4 as [layer 3][kernel column 3][ci chunk 4][(kernel row, co) 96][8] bf16
#define AZ_TRUNK_ROWS_MIN_G 1024        // smaller batches keep the first kernel (smaller groups spread over more SMs)

bool az_trunk_rows_enabled();
int az_trunk_rows_launch(const uint64_t* own, const uint64_t* opp, int64_t G, const void* wblob, void* features,
                     int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream);

// k_net_trunk_rows2<rows, cols> (trunk_rows2_kernels.cu): the same formulation with two CTAs per SM, for the 8 x 8 image
// (14 boards per group) and Connect4Net's 7 x 6 image (18 boards per group)
bool az_trunk_rows2_enabled();
int az_trunk_rows2_boards_per_group(int rows);
int az_trunk_rows2_launch(const uint64_t* own, const uint64_t* opp, int rows, int cols, int64_t G, const void* wblob, void* features,
                          int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream);
