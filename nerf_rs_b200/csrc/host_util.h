// host_util.h — host-side pieces of the boundary: weight files, camera JSON, PPM.
#pragma once
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/nerf_b200.h"

namespace nerf {

void set_error(const std::string& msg);
const char* get_error();

struct HostTensor {
  std::vector<int64_t> dims;
  std::vector<float> data;
};
using TensorMap = std::map<std::string, HostTensor>;

// load_shapes + load_tensor (src/lib.rs:35-42,63-74): every tensor listed in <dir>/shapes.txt.
int load_tensor_dir(const std::string& dir, TensorMap* out);
// The 24 names load_network_from_dir takes (src/lib.rs:133-169) with the dims of shapes.txt.
int validate_network(const TensorMap& t, const char* which);

// f32 Vec3 helpers with the reference's operation order (src/vec3.rs:19-34).
void normalize3(const float in[3], float out[3]);
void cross3(const float a[3], const float b[3], float out[3]);

}  // namespace nerf
