// Host driver of the clip-rescue phase + brb_ksw_extend (included at the end of engine.cu).
namespace {
int run_rescue(brb_ctx* c, const brb_params& p) {
  (void)c; (void)p;
  return fail(BRB_E_UNSUPPORTED, "clip rescue not built yet");
}
}  // namespace

extern "C" int brb_ksw_extend(brb_ctx* c, uint32_t n, const uint64_t* q_off, const uint8_t* q, const uint64_t* t_off, const uint8_t* t,
                              int32_t* max, int32_t* score, int32_t* max_q, int32_t* max_t, uint8_t* zdropped, uint32_t cigar_cap,
                              uint32_t* n_cigar, uint32_t* cigar_out) {
  return fail(BRB_E_UNSUPPORTED, "ksw not built yet");
}
