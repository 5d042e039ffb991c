// codestream.cpp -- marker segments of a JPEG 2000 codestream (host only, no CUDA): main header,
// tile-part headers and the index of tile-part bodies the packet parser (t2_parse.cpp) walks.
// T2 on our side of the plugin.
//
// Follows ISO/IEC 15444-1 Annex A as the reference reads it (paths under
// /root/reference/src/lib/jp2):
//   SIZ                 codestream/markers/SIZMarker.cpp:28-260
//   COD / COC           codestream/CodeStreamDecompress.cpp:1263-1475, SPcod :1709-1860
//   QCD / QCC scoping   codestream/Quantizer.cpp:64-98, :160-345 (main QCD < main QCC < tile QCD <
//                       tile QCC; scalar-derived exponents :330-343)
//   RGN                 codestream/CodeStreamDecompress.cpp:1477-1525
//   POC                 codestream/CodeStreamDecompress.cpp:1148-1245
//   SOT / tile parts    codestream/markers/SOTMarker.cpp:60-240
// What this parser refuses (loudly; the caller fails the frame): sub-sampled components, more than
// GRKB_MAX_COMPS components, component-specific COC geometry, packed packet headers (PPM / PPT).
#include "codestream.h"
#include <string.h>

namespace grkb_cs
{
namespace
{
enum : uint16_t
{
	SOC = 0xFF4F, CAP = 0xFF50, SIZ = 0xFF51, COD = 0xFF52, COC = 0xFF53, TLM = 0xFF55, PLM = 0xFF57,
	PLT = 0xFF58, CPF = 0xFF59, QCD = 0xFF5C, QCC = 0xFF5D, RGN = 0xFF5E, POC = 0xFF5F, PPM = 0xFF60,
	PPT = 0xFF61, CRG = 0xFF63, COM = 0xFF64, SOT = 0xFF90, SOD = 0xFF93, EOC = 0xFFD9
};

inline uint32_t be16(const uint8_t* p) { return ((uint32_t)p[0] << 8) | p[1]; }
inline uint32_t be32(const uint8_t* p)
{
	return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

bool fail(std::string& err, const char* what)
{
	err = what;
	return false;
}

// locate the contiguous codestream inside a JP2 file (ISO/IEC 15444-1 Annex I: 'jp2c' box)
bool find_codestream(const uint8_t* d, size_t len, size_t& off, size_t& cs_len)
{
	if(len >= 2 && be16(d) == SOC)
	{
		off = 0;
		cs_len = len;
		return true;
	}
	static const uint8_t sig[12] = {0, 0, 0, 12, 'j', 'P', ' ', ' ', 0x0D, 0x0A, 0x87, 0x0A};
	if(len < 12 || memcmp(d, sig, 12) != 0)
		return false;
	size_t p = 0;
	while(p + 8 <= len)
	{
		uint64_t lbox = be32(d + p);
		const uint32_t tbox = be32(d + p + 4);
		size_t hdr = 8;
		if(lbox == 1)
		{
			if(p + 16 > len)
				return false;
			lbox = ((uint64_t)be32(d + p + 8) << 32) | be32(d + p + 12);
			hdr = 16;
		}
		else if(lbox == 0)
			lbox = len - p;
		if(lbox < hdr || lbox > len - p)
			return false;
		if(tbox == 0x6A703263u) // 'jp2c'
		{
			off = p + hdr;
			cs_len = (size_t)lbox - hdr;
			return cs_len >= 2 && be16(d + off) == SOC;
		}
		p += (size_t)lbox;
	}
	return false;
}

// SPcod / SPcoc
bool read_spcod(const uint8_t* p, size_t n, bool precincts, Coding& c, std::string& err)
{
	if(n < 5)
		return fail(err, "COD/COC: segment too short");
	if(p[0] > 32)
		return fail(err, "COD/COC: more than 32 decomposition levels");
	c.numres = (uint8_t)(p[0] + 1);
	if(p[1] > 8 || p[2] > 8 || p[1] + p[2] > 8)
		return fail(err, "COD/COC: illegal code-block size");
	c.cblkw_exp = (uint8_t)(p[1] + 2);
	c.cblkh_exp = (uint8_t)(p[2] + 2);
	c.cblk_sty = p[3];
	if((c.cblk_sty & 0x40) && c.cblk_sty != 0x40)
		return fail(err, "COD/COC: HT code-block style with Part-1 mode switches (mixed mode) is not supported");
	if(c.cblk_sty & 0x80)
		return fail(err, "COD/COC: unrecognised code-block style");
	if(p[4] > 1)
		return fail(err, "COD/COC: unknown wavelet transform");
	c.reversible = p[4];
	for(uint32_t r = 0; r < MAX_RES; ++r)
		c.pw[r] = c.ph[r] = 15;
	if(precincts)
	{
		if(n < 5u + c.numres)
			return fail(err, "COD/COC: precinct sizes missing");
		for(uint32_t r = 0; r < c.numres; ++r)
		{
			c.pw[r] = p[5 + r] & 0xF;
			c.ph[r] = p[5 + r] >> 4;
			if(r > 0 && (c.pw[r] == 0 || c.ph[r] == 0))
				return fail(err, "COD/COC: precinct size 1 above the lowest resolution");
		}
	}
	return true;
}

bool read_sqcd(const uint8_t* p, size_t n, Quant& q, std::string& err)
{
	if(n < 1)
		return fail(err, "QCD/QCC: segment too short");
	q.qntsty = p[0] & 0x1F;
	q.numgbits = p[0] >> 5;
	if(q.qntsty > 2)
		return fail(err, "QCD/QCC: unknown quantisation style");
	memset(q.expn, 0, sizeof(q.expn));
	memset(q.mant, 0, sizeof(q.mant));
	++p;
	--n;
	if(q.qntsty == 0)
	{
		for(size_t b = 0; b < n && b < MAX_BANDS; ++b)
			q.expn[b] = p[b] >> 3;
	}
	else
	{
		const size_t nb = q.qntsty == 1 ? 1 : n / 2;
		if(n < 2)
			return fail(err, "QCD/QCC: step sizes missing");
		for(size_t b = 0; b < nb && b < MAX_BANDS; ++b)
		{
			const uint32_t v = be16(p + 2 * b);
			q.expn[b] = (uint8_t)(v >> 11);
			q.mant[b] = (uint16_t)(v & 0x7FF);
		}
		if(q.qntsty == 1) // scalar derived (Quantizer.cpp:330-343)
			for(uint32_t b = 1; b < MAX_BANDS; ++b)
			{
				const uint32_t k = (b - 1) / 3;
				q.expn[b] = q.expn[0] > k ? (uint8_t)(q.expn[0] - k) : 0;
				q.mant[b] = q.mant[0];
			}
	}
	return true;
}

// one marker segment of a main or tile-part header applied to `h`
bool apply_marker(Codestream& cs, TileHeader& h, bool tile, uint16_t id, const uint8_t* p, size_t n,
				  std::string& err)
{
	const uint32_t nc = cs.numcomps;
	const size_t croom = nc <= 256 ? 1 : 2;
	switch(id)
	{
		case COD: {
			if(n < 5)
				return fail(err, "COD: segment too short");
			Coding c;
			const uint8_t scod = p[0];
			if(!read_spcod(p + 5, n - 5, (scod & 1) != 0, c, err))
				return false;
			c.level = tile ? 3 : 1;
			h.csty = scod;
			h.prog = p[1];
			h.numlayers = (uint16_t)be16(p + 2);
			h.mct = p[4];
			if(h.prog > 4)
				return fail(err, "COD: unknown progression order");
			if(h.numlayers == 0)
				return fail(err, "COD: zero layers");
			for(uint32_t i = 0; i < nc; ++i)
				if(h.cod[i].level <= c.level || (tile && h.cod[i].level == 2))
					h.cod[i] = c;
			return true;
		}
		case COC: {
			if(n < croom + 1)
				return fail(err, "COC: segment too short");
			const uint32_t comp = croom == 1 ? p[0] : be16(p);
			if(comp >= nc)
				return fail(err, "COC: component out of range");
			Coding c;
			if(!read_spcod(p + croom + 1, n - croom - 1, (p[croom] & 1) != 0, c, err))
				return false;
			c.level = tile ? 4 : 2;
			if(h.cod[comp].level <= c.level)
				h.cod[comp] = c;
			return true;
		}
		case QCD: {
			Quant q;
			if(!read_sqcd(p, n, q, err))
				return false;
			q.level = tile ? 3 : 1;
			for(uint32_t i = 0; i < nc; ++i)
				if(h.q[i].level <= q.level || (tile && h.q[i].level == 2))
					h.q[i] = q;
			return true;
		}
		case QCC: {
			if(n < croom + 1)
				return fail(err, "QCC: segment too short");
			const uint32_t comp = croom == 1 ? p[0] : be16(p);
			if(comp >= nc)
				return fail(err, "QCC: component out of range");
			Quant q;
			if(!read_sqcd(p + croom, n - croom, q, err))
				return false;
			q.level = tile ? 4 : 2;
			if(h.q[comp].level <= q.level)
				h.q[comp] = q;
			return true;
		}
		case RGN: {
			if(n < croom + 2)
				return fail(err, "RGN: segment too short");
			const uint32_t comp = croom == 1 ? p[0] : be16(p);
			if(comp >= nc)
				return fail(err, "RGN: component out of range");
			if(p[croom] != 0)
				return fail(err, "RGN: unknown ROI style");
			if(p[croom + 1] >= 32)
				return fail(err, "RGN: unsupported ROI shift");
			h.roishift[comp] = p[croom + 1];
			return true;
		}
		case POC: {
			const size_t rec = 5 + 2 * croom;
			if(n == 0 || n % rec != 0)
				return fail(err, "POC: bad segment length");
			for(size_t i = 0; i < n / rec; ++i)
			{
				const uint8_t* q = p + i * rec;
				Poc r;
				r.resS = q[0];
				r.compS = (uint16_t)(croom == 1 ? q[1] : be16(q + 1));
				r.layE = (uint16_t)be16(q + 1 + croom);
				r.resE = q[3 + croom];
				r.compE = (uint16_t)(croom == 1 ? q[4 + croom] : be16(q + 4 + croom));
				if(croom == 1 && r.compE == 0)
					r.compE = 256;
				r.prog = q[4 + 2 * croom];
				if(r.prog > 4)
					return fail(err, "POC: unknown progression order");
				if(r.compE > nc)
					r.compE = (uint16_t)nc;
				if(r.resE <= r.resS || r.compE <= r.compS)
					return fail(err, "POC: empty resolution or component range");
				if(h.pocs.size() >= 32)
					return fail(err, "POC: too many progression order changes");
				h.pocs.push_back(r);
			}
			return true;
		}
		case PPM:
		case PPT:
			return fail(err, "packed packet headers (PPM / PPT) are not supported");
		default:
			return true; // TLM, PLM, PLT, CRG, COM, CAP, CPF, ...: nothing the tile decode needs
	}
}
} // namespace

bool open(const uint8_t* data, size_t len, Codestream& cs, std::string& err)
{
	cs = Codestream();
	cs.data = data;
	cs.len = len;
	size_t off = 0, cs_len = 0;
	if(!data || !find_codestream(data, len, off, cs_len))
		return fail(err, "no JPEG 2000 codestream found (neither SOC nor a JP2 file with a jp2c box)");
	cs.cs_off = off;
	const uint8_t* d = data + off;
	size_t p = 2;
	// ---- main header
	if(p + 4 > cs_len || be16(d + p) != SIZ)
		return fail(err, "SIZ does not follow SOC");
	bool have_cod = false, have_qcd = false;
	while(true)
	{
		if(p + 4 > cs_len)
			return fail(err, "main header truncated");
		const uint16_t id = (uint16_t)be16(d + p);
		if(id == SOT)
			break;
		if(id < 0xFF00)
			return fail(err, "main header: expected a marker");
		const size_t L = be16(d + p + 2);
		if(L < 2 || p + 2 + L > cs_len)
			return fail(err, "main header: marker segment runs past the end");
		const uint8_t* body = d + p + 4;
		const size_t n = L - 2;
		if(id == SIZ)
		{
			if(n < 36)
				return fail(err, "SIZ: segment too short");
			cs.rsiz = (uint16_t)be16(body);
			cs.x1 = be32(body + 2);
			cs.y1 = be32(body + 6);
			cs.x0 = be32(body + 10);
			cs.y0 = be32(body + 14);
			cs.tw = be32(body + 18);
			cs.th = be32(body + 22);
			cs.tx0 = be32(body + 26);
			cs.ty0 = be32(body + 30);
			const uint32_t nc = be16(body + 34);
			if(nc == 0 || n < 36 + 3 * (size_t)nc)
				return fail(err, "SIZ: component descriptions missing");
			if(nc > GRKB_MAX_COMPS)
				return fail(err, "more components than GRKB_MAX_COMPS");
			cs.numcomps = nc;
			for(uint32_t c = 0; c < nc; ++c)
			{
				const uint8_t* s = body + 36 + 3 * c;
				cs.comps[c].sgnd = s[0] >> 7;
				cs.comps[c].prec = (uint8_t)((s[0] & 0x7F) + 1);
				cs.comps[c].dx = s[1];
				cs.comps[c].dy = s[2];
				if(s[1] != 1 || s[2] != 1)
					return fail(err, "sub-sampled components are not supported");
				if(cs.comps[c].prec > 31)
					return fail(err, "component precision above 31 bits");
			}
			if(cs.x1 <= cs.x0 || cs.y1 <= cs.y0 || cs.tw == 0 || cs.th == 0 || cs.tx0 > cs.x0 || cs.ty0 > cs.y0 ||
			   (uint64_t)cs.tx0 + cs.tw <= cs.x0 || (uint64_t)cs.ty0 + cs.th <= cs.y0)
				return fail(err, "SIZ: inconsistent image / tile geometry");
			cs.ntx = (uint32_t)(((uint64_t)cs.x1 - cs.tx0 + cs.tw - 1) / cs.tw);
			cs.nty = (uint32_t)(((uint64_t)cs.y1 - cs.ty0 + cs.th - 1) / cs.th);
			if((uint64_t)cs.ntx * cs.nty > 65535)
				return fail(err, "SIZ: more than 65535 tiles");
		}
		else
		{
			if(cs.numcomps == 0)
				return fail(err, "marker before SIZ");
			if(id == CAP)
				cs.cap = true;
			if(!apply_marker(cs, cs.main, false, id, body, n, err))
				return false;
			have_cod |= id == COD;
			have_qcd |= id == QCD;
		}
		p += 2 + L;
	}
	if(!have_cod || !have_qcd)
		return fail(err, "main header without COD or QCD");
	cs.tiles.assign((size_t)cs.ntx * cs.nty, Tile());
	// ---- tile parts
	while(p + 2 <= cs_len)
	{
		const uint16_t id = (uint16_t)be16(d + p);
		if(id == EOC)
			break;
		if(id != SOT)
			return fail(err, "expected SOT or EOC between tile parts");
		if(p + 12 > cs_len || be16(d + p + 2) != 10)
		{
			cs.truncated = true;
			break;
		}
		const uint32_t isot = be16(d + p + 4);
		uint64_t psot = be32(d + p + 6);
		if(isot >= cs.tiles.size())
			return fail(err, "SOT: tile index out of range");
		const size_t start = p;
		size_t end;
		if(psot == 0)
		{
			end = cs_len;
			if(end >= 2 && be16(d + end - 2) == EOC)
				end -= 2;
		}
		else
		{
			if(psot < 14)
				return fail(err, "SOT: tile-part length too small");
			end = start + (size_t)psot;
			if(end > cs_len)
			{
				end = cs_len;
				cs.truncated = true;
			}
		}
		Tile& t = cs.tiles[isot];
		if(!t.seen)
		{
			t.hdr = cs.main;
			t.seen = true;
		}
		p += 12;
		bool sod = false;
		while(p + 2 <= end)
		{
			const uint16_t m = (uint16_t)be16(d + p);
			if(m == SOD)
			{
				p += 2;
				sod = true;
				break;
			}
			if(m < 0xFF00 || p + 4 > end)
				return fail(err, "tile-part header: expected a marker");
			const size_t L = be16(d + p + 2);
			if(L < 2 || p + 2 + L > end)
				return fail(err, "tile-part header: marker segment runs past the tile part");
			if(!apply_marker(cs, t.hdr, true, m, d + p + 4, L - 2, err))
				return false;
			p += 2 + L;
		}
		if(!sod)
		{
			if(cs.truncated)
				break;
			return fail(err, "tile part without SOD");
		}
		if(end > p)
			t.parts.push_back(Range{off + p, end - p});
		p = end;
		if(cs.truncated)
			break;
	}
	return true;
}
} // namespace grkb_cs
