/* TEST INFRASTRUCTURE ONLY.  Compiles the DEVICE decoder's core (deepore_b200/csrc/inflate_core.h, the code that
 * bgzf_inflate_kernel runs per thread) for the host, so that the CPU test suite can check it against zlib (the library
 * the reference's htslib calls in dep/htslib/bgzf.c:456-499) without a GPU.  Never loaded by the product. */
#include "../deepore_b200/csrc/inflate_core.h"

int dpo_inflate(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t out_cap, uint32_t *out_len)
{
    uint32_t scratch32[(DPI_SCRATCH_U16 + 3) / 2];            /* 4-byte aligned: the lookup tables hold 32-bit entries */
    uint16_t *scratch = (uint16_t *)scratch32;
    return dpi_inflate(in, in_len, out, out_cap, scratch, out_len);
}
