/* oracle/ref_rand.c -- TEST INFRASTRUCTURE ONLY.
 * The reference draws the FracMinHash parameters (x, y) from rand() seeded with time()
 * (tools/hash_function_generator.h:91-117), so two runs never build the same index.  Linking this file into
 * oracle/_ref/bucketmap* replaces libc's rand()/srand(): the first two draws come from the environment variable
 * BM_RAND="a,b" (x = a % (p-1) + 1, y = b % p), later draws from a fixed LCG.  The reference sources stay unmodified. */
#include <stdlib.h>
#include <string.h>
static unsigned long long state = 12345;
static int served = 0;
void srand(unsigned int seed) { (void)seed; }
int rand(void) {
    const char* e = getenv("BM_RAND");
    if (e && served < 2) {
        long a = 0, b = 0;
        const char* c = strchr(e, ',');
        a = atol(e);
        if (c) b = atol(c + 1);
        return (int)(served++ == 0 ? a : b);
    }
    state = state * 6364136223846793005ULL + 1442695040888963407ULL;
    return (int)((state >> 33) & 0x7fffffff);
}
