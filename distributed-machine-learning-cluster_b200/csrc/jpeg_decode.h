// JPEG decode on the GPU in front of the resize kernels, so that a `predict` that starts from file
// contents (reference src/services.rs:484-495: read_dir -> load_image_and_resize -> forward) no longer
// decodes one image at a time on the calling thread.  The decoder is NVIDIA's nvJPEG, loaded with
// dlopen at first use: libb200infer.so itself keeps linking nothing but the static CUDA runtime, and
// every entry point that needs the decoder reports B2I_ENODEV-style errors when the library is absent.
// Backends: the NVJPG hardware engines where the device exposes them (baseline JPEG, batched), else the hybrid
// decoder (entropy decoding on a pool of host threads, one nvJPEG state each; IDCT and colour conversion on the
// GPU); GPU-assisted Huffman decode on request (B2I_JPEG_BACKEND=gpu; slower than the host threads here).
//
// Decoders differ by +-1-2 LSB per pixel (stb in the reference, libjpeg-turbo under PIL, nvJPEG here;
// SURVEY.md 8c R2), so parity of this stage is defined from the decoded pixels on: the tests read them
// back with b2i_jpeg_decode_u8 and hand THOSE to the CPU restatement.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include <string>

namespace b2i {

struct JpegDecoder;

// Decoder bound to the current device; nullptr + *err when nvJPEG cannot be loaded or initialised.
JpegDecoder* jpeg_decoder_create(std::string* err);
void jpeg_decoder_destroy(JpegDecoder* d);
const char* jpeg_decoder_backend(const JpegDecoder* d);

// Header parse only (host): h[i], w[i] and status[i] = 0, or status[i] = B2I_EDECODE when the stream is
// not a JPEG nvJPEG can decode (null / empty / corrupt header / unsupported coding such as CMYK).
void jpeg_probe(JpegDecoder* d, const uint8_t* const* data, const size_t* len, int n, int* h, int* w,
                int* status);

// Decode every image whose status is 0 to interleaved RGB u8 [h][w][3] at d_out[i] (device memory, row
// pitch w * 3), enqueuing the GPU work on `st`.  An image whose bitstream turns out to be damaged gets
// status B2I_EDECODE; the others are still decoded.  Returns false only for failures of the decoder
// itself.  The call may block on host-side entropy decoding, never on the GPU work.
bool jpeg_decode_batch(JpegDecoder* d, const uint8_t* const* data, const size_t* len, int n,
                       uint8_t* const* d_out, const int* h, const int* w, int* status, cudaStream_t st,
                       std::string* err);

}  // namespace b2i
