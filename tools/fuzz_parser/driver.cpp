// ASan/UBSan driver for the host parser: every file named on the command line is opened and every tile parsed
#include "codestream.h"
#include <stdarg.h>
#include <string.h>
#include <stdio.h>
#include <vector>
#include <string>
void grkb_set_error(const char* fmt, ...) { (void)fmt; }
extern "C" const char* grkb_last_error(void) { return ""; }
int main(int argc, char** argv)
{
	size_t ok = 0, bad = 0;
	for(int i = 1; i < argc; ++i)
	{
		FILE* f = fopen(argv[i], "rb");
		if(!f) continue;
		std::vector<uint8_t> b;
		uint8_t tmp[65536]; size_t n;
		while((n = fread(tmp, 1, sizeof tmp, f)) > 0) b.insert(b.end(), tmp, tmp + n);
		fclose(f);
		// exact-size heap copy so that any over-read is caught
		uint8_t* d = new uint8_t[b.size() ? b.size() : 1];
		memcpy(d, b.data(), b.size());
		grkb_cs::Codestream cs; std::string err;
		if(grkb_cs::open(d, b.size(), cs, err))
		{
			const uint32_t nt = cs.ntx * cs.nty;
			for(uint32_t t = 0; t < nt && t < 8; ++t)
				for(uint32_t v = 0; v < 3; ++v)
				{
					grkb_cs::TileTables tt;
					uint32_t win[4] = {3, 2, 40, 30};
					if(grkb_cs::parse_tile(cs, t, v == 1 ? 1 : 0, v == 2 ? 1 : 0, tt, err, v == 2 ? win : nullptr))
					{
						++ok;
						// what the device would be given must lie inside what it is given
						const size_t pool = tt.zero_copy ? tt.src_len : tt.pool.size();
						for(const auto& c : tt.blocks)
						{
							if((size_t)c.data_off + c.data_len > pool + 16) { fprintf(stderr, "block bytes out of range in %s\n", argv[i]); return 3; }
							if((size_t)c.seg_first + c.numsegs > tt.segs.size()) { fprintf(stderr, "segments out of range in %s\n", argv[i]); return 3; }
							size_t sl = 0;
							for(uint32_t s = 0; s < c.numsegs; ++s) sl += tt.segs[c.seg_first + s].len;
							if(sl > c.data_len) { fprintf(stderr, "segment lengths exceed the block's bytes in %s\n", argv[i]); return 3; }
						}
						if(tt.zero_copy && tt.src_off + tt.src_len > b.size()) { fprintf(stderr, "source range out of the file in %s\n", argv[i]); return 3; }
					}
					else ++bad;
				}
		}
		else ++bad;
		delete[] d;
	}
	printf("parsed %zu refused %zu\n", ok, bad);
	return 0;
}
