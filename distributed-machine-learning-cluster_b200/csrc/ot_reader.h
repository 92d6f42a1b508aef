// Reader for the `.ot` weight archives the reference loads with `vs.load(...)`
// (reference src/services.rs:516,522; format dissected in SURVEY.md Appendix A): a ZIP container
// whose `<stem>/data.pkl` is a protocol-2 pickle naming fp32 tensors stored raw in
// `<stem>/data/<key>`.  No libtorch: a central-directory walk plus a minimal pickle VM.
#pragma once
#include <map>
#include <string>
#include <vector>
#include <stdint.h>

namespace b2i {

struct HostTensor {
  std::vector<int64_t> shape;
  std::vector<float> data;  // contiguous, row-major, converted to fp32
  int64_t numel() const {
    int64_t n = 1;
    for (int64_t d : shape) n *= d;
    return n;
  }
};

// Names are normalised to '.'-separated paths ("layer1.0.conv1.weight"); the archive may use '|'
// (what tch writes) or '.'.  Returns false and fills `err` on any malformed input.
bool read_ot_archive(const std::string& path, std::map<std::string, HostTensor>* out,
                     std::string* err);

}  // namespace b2i
