// Batched beam search (Decoder.caption, decoder.py:160-269).  Stub until the training path is verified.
#pragma once
#include "sat_decoder.cuh"

template <typename T> struct BeamWs { int64_t total; };
static inline int beam_check_args(const sat_beam_args* a) {
    SAT_REQUIRE(a != nullptr, "null args");
    SAT_REQUIRE(a->struct_bytes == (int32_t)sizeof(sat_beam_args), "sat_beam_args size mismatch");
    return SAT_OK;
}
template <typename T> int64_t beam_carve(const sat_beam_args&, BeamWs<T>& w, uint8_t*) { w.total = 0; return 0; }
template <typename T> int beam_search_t(const sat_beam_args&) { sat_set_error("beam search not built yet"); return SAT_ERR_UNSUPPORTED; }
