// weight_cache.h — pre-swizzled weight cache file (weight_cache.cu).
#pragma once
#include <string>
#include <vector>

#include "mlp_tc.h"

namespace nerf {

struct CachedWeights {
  TensorMap nets[2];                              // coarse, fine: the f32 tensors of shapes.txt
  std::vector<uint8_t> images[2][kNumTcFormats];  // chunk streams per operand format
};

int serialize_weight_cache(const TensorMap nets[2], std::vector<uint8_t>* out);
int parse_weight_cache(const uint8_t* data, size_t size, CachedWeights* out);
int write_file(const std::string& path, const std::vector<uint8_t>& b);
int read_file(const std::string& path, std::vector<uint8_t>* b);

}  // namespace nerf
