/* oracle/ref_bits.h — MSB-first bit writer used by the CPU oracle. TEST INFRASTRUCTURE ONLY. */
#ifndef REF_BITS_H
#define REF_BITS_H
#include <stdint.h>
void ref_put_bits(uint8_t *out, int *bitpos, int n, uint32_t v);
void ref_put_ue(uint8_t *out, int *bitpos, unsigned v);
void ref_put_se(uint8_t *out, int *bitpos, int v);
#endif
