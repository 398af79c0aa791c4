// Host check of image_codecs_b200/csrc/jpeg_unstuff.h: the word-at-a-time unstuffing (what k_unstuff runs per
// lane) against the plain byte loop with the reference's rules, on random streams dense in FF / 00 / markers.
// Build: g++ -O1 -I image_codecs_b200/csrc tests/native/unstuff_check.cpp -o unstuff_check ; exit code 0 = all equal.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "jpeg_unstuff.h"

static uint32_t rng_state = 12345u;
static uint32_t rnd()
{
	rng_state ^= rng_state << 13;
	rng_state ^= rng_state >> 17;
	rng_state ^= rng_state << 5;
	return rng_state;
}

static uint32_t byte_loop(const uint8_t *base, uint32_t len, uint32_t start, uint32_t *out)
{
	UnstuffOut o;
	uint32_t pos = start;
	o.wp = out;
	o.pend = 0u;
	o.nacc = 0u;
	o.words = 0u;
	while (unstuff_byte(base, pos, len, o))
	{
	}
	if (o.nacc)
	{
		*o.wp++ = o.pend;
		++o.words;
	}
	return o.words;
}

int main(int argc, char **argv)
{
	const int cases = argc > 1 ? atoi(argv[1]) : 20000;
	std::vector<uint8_t> buf(512 + 64);
	std::vector<uint32_t> a(256), b(256);
	for (int c = 0; c < cases; ++c)
	{
		const uint32_t len = 1 + rnd() % 300;
		const int flavour = (int)(rnd() % 4);
		for (uint32_t i = 0; i < buf.size(); ++i)
		{
			const uint32_t r = rnd() % 100;
			uint8_t v = (uint8_t)rnd();
			if (flavour == 0) /* mostly stuffed pairs */
				v = r < 30 ? 0xff : r < 70 ? 0x00 : v;
			else if (flavour == 1) /* realistic: rare FF */
				v = r < 2 ? 0xff : r < 6 ? 0x00 : v;
			else if (flavour == 2) /* fill runs and markers */
				v = r < 50 ? 0xff : r < 70 ? 0x00 : r < 75 ? 0xd0 + (uint8_t)(rnd() % 8) : v;
			buf[i] = v; /* bytes past len are garbage on purpose: they must not matter */
		}
		if (flavour == 1) /* no accidental marker: every FF gets its 00, except one real marker somewhere */
		{
			for (uint32_t i = 0; i + 1 < len; ++i)
				if (buf[i] == 0xff)
					buf[i + 1] = 0x00;
			if (len > 4 && (rnd() & 1))
			{
				const uint32_t m = rnd() % (len - 1);
				buf[m] = 0xff;
				buf[m + 1] = 0xd3;
			}
		}
		const uint32_t start = rnd() % len;
		std::fill(a.begin(), a.end(), 0u);
		std::fill(b.begin(), b.end(), 0u);
		const uint32_t na = byte_loop(buf.data(), len, start, a.data());
		const uint32_t nb = unstuff_segment_scalar(buf.data(), len, start, b.data());
		if (na != nb || memcmp(a.data(), b.data(), sizeof(uint32_t) * na) != 0)
		{
			printf("MISMATCH case %d flavour %d len %u start %u words %u vs %u\n", c, flavour, len, start, na, nb);
			for (uint32_t i = start; i < len; ++i)
				printf("%02x ", buf[i]);
			printf("\n");
			return 1;
		}
	}
	printf("ok %d cases\n", cases);
	return 0;
}
