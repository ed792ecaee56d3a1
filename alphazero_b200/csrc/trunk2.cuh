// k_net_trunk2 (trunk2_kernels.cu): the shuffle-free convolution trunk for 8 x 8 boards; launched by az_net_trunk.
#pragma once
#include <stdint.h>

#define AZ_TRUNK2_BOARDS_PER_GROUP 14
#define AZ_TRUNK2_MIN_G 1024        // smaller batches keep the first kernel (smaller groups spread over more SMs)

bool az_trunk2_enabled();
int az_trunk2_launch(const uint64_t* own, const uint64_t* opp, int64_t G, const void* wblob, void* features,
                     int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream);
