/* TEST INFRASTRUCTURE ONLY (oracle/) -- see orb_oracle.h.
 * CPU restatement of the Hamming core of ORBmatcher (src/ORBmatcher.cc). */
#include "orb_oracle.h"
#include <stdlib.h>
#include <string.h>

/* DescriptorDistance: src/ORBmatcher.cc:2123-2143 (8 x 32-bit bit-hack popcount) */
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t x, y;
        memcpy(&x, a + 4 * i, 4); memcpy(&y, b + 4 * i, 4);
        uint32_t v = x ^ y;
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
    }
    return dist;
}
