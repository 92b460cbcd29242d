/*
 * oracle/hbtseries_api.h -- CPU checkers for the series reductions dynolog applies to a
 * metric's samples before logging them (the sink side).
 * TEST INFRASTRUCTURE ONLY, like hbtref_api.h.
 *
 *   hbtref_series_reduce     oracle/_ref/libhbtref.so: the REAL
 *                            facebook::dynolog::MetricSeries<uint64_t>
 *                            (dynolog/src/metric_frame/MetricSeries.h:198-275):
 *                            avg<double>, avg<uint64_t>, percentile, diff, rate<double>,
 *                            rate<uint64_t>, filled with addSample (:153-163)
 *   hbtoracle_series_reduce  oracle/libhbtoracle.so: plain-C restatement (hbt_oracle.c)
 *
 * series: n_series rows of len u64 samples, row-major (row s at series + s * len).
 * percentage in [0, 1]; period_ns / duration_ns are std::chrono::steady_clock::duration ticks.
 * Every output pointer may be NULL; each has n_series entries.
 */
#ifndef HBTSERIES_API_H
#define HBTSERIES_API_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HBTSERIES_DECL(prefix)                                                               \
  int prefix##_series_reduce(const uint64_t* series, uint64_t n_series, uint64_t len,        \
      float percentage, int64_t period_ns, int64_t duration_ns, double* avg_f64,             \
      uint64_t* avg_u64, uint64_t* percentile, uint64_t* diff, double* rate_f64,             \
      uint64_t* rate_u64);

HBTSERIES_DECL(hbtref)
HBTSERIES_DECL(hbtoracle)

#ifdef __cplusplus
}
#endif
#endif
