// tcgen05/TMEM extend kernel (included by extend.cu inside namespace b200).
// Placeholder until the tensor-core path lands: reports "not supported" so the
// dispatcher uses the generic kernel.
namespace {

bool extend_fast_supported(int, int, int, int, int) { return false; }
int extend_fast_block_q(int) { return 64; }
bool extend_fast_layout_ok(const ExtendParams&, int) { return false; }
int launch_extend_fast(const ExtendParams&, const b200_extend_args&, cudaStream_t) {
  set_error("extend: fast kernel not built");
  return B200_EINVAL;
}

}  // namespace
