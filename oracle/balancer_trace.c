/*
 * oracle/balancer_trace.c — drives the REFERENCE's own balancer (compiled from /root/reference/src/core/*.c where
 * the sources lie; nothing is copied) at its 20 ms tick (/root/reference/src/core/bitrate_control.h:35) to produce
 * the bitrate schedule of BASELINE config 2 ("balancer bitrate steps 2->12 Mbps"): good network, one 2 s congestion
 * burst at t = 10 s -> emergency drop to min, then the +100..300 kbps / 520 ms ladder back to max
 * (/root/reference/src/core/bitrate_control.c:176-206).  TEST INFRASTRUCTURE ONLY: output is committed as
 * tests/golden/balancer_2_12mbps.csv and replayed into cc_enc_set_bitrate by the rate-control tests and bench.py.
 *
 *   balancer_trace [algo=adaptive] [min_kbps=2000] [max_kbps=12000] [seconds=40]  ->  "t_ms,bitrate_bps" on change
 */
#include "balancer.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char **argv) {
    const char *name = argc > 1 ? argv[1] : "adaptive";
    int min_kbps = argc > 2 ? atoi(argv[2]) : 2000, max_kbps = argc > 3 ? atoi(argv[3]) : 12000, seconds = argc > 4 ? atoi(argv[4]) : 40;
    const BalancerAlgorithm *a = balancer_find(name);
    if (!a) { fprintf(stderr, "no balancer '%s'\n", name); return 1; }
    BalancerConfig c;
    memset(&c, 0, sizeof c);
    c.min_bitrate = min_kbps * 1000; c.max_bitrate = max_kbps * 1000; c.srt_latency = 2000; c.srt_pkt_size = 1316;
    c.adaptive_incr_step = 30000; c.adaptive_decr_step = 100000; c.adaptive_incr_interval = 500; c.adaptive_decr_interval = 200; /* config.c:27-43 defaults */
    c.aimd_incr_step = 50000; c.aimd_decr_mult = 0.75; c.aimd_incr_interval = 500; c.aimd_decr_interval = 200;
    void *st = a->init(&c);
    if (!st) return 2;
    int last = -1;
    printf("t_ms,bitrate_bps\n");
    for (int t = 0; t <= seconds * 1000; t += 20) {
        BalancerInput in;
        memset(&in, 0, sizeof in);
        const int burst = t >= 10000 && t < 12000;
        in.buffer_size = burst ? 300 : 10; in.rtt = burst ? 700.0 : 30.0; in.send_rate_mbps = 5.0; in.timestamp = (uint64_t)t + 1000;
        BalancerOutput o = a->step(st, &in);
        if (o.new_bitrate != last) { printf("%d,%d\n", t, o.new_bitrate); last = o.new_bitrate; }
    }
    a->cleanup(st);
    return 0;
}
