// k_net_trunk3 (trunk3_kernels.cu): the convolution trunk for 8 x 8 boards with one TMEM accumulator per output tile;
// launched by az_net_trunk.
#pragma once
#include <stdint.h>

#define AZ_TRUNK3_BOARDS_PER_GROUP 14
#define AZ_TRUNK3_MIN_G 1024        // smaller batches keep the first kernel (smaller groups spread over more SMs)

bool az_trunk3_enabled();
int az_trunk3_launch(const uint64_t* own, const uint64_t* opp, int64_t G, const void* wblob, void* features,
                     int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream);
