// gs2d_jpeg.cpp — target ingest, SURVEY.md §8 f2: JPEG -> RGBA8 with the pixel values the reference sees.
//
// The reference loads its target with prlib's Image2DRGBA8::load (prp.cpp:179-193) = stb_image v2.23's stbi_load with
// 4 requested channels (libs/prlib/libs/stb/stb_image.h). JPEG decoding is only specified up to IDCT accuracy, chroma
// up-sampling filter and colour-matrix rounding, so "the same picture" means reproducing stb's choices:
//   * IDCT: the integer "islow" transform with 12-bit constants, 2 extra bits kept after the column pass, +128 level
//     shift folded into the row pass rounding term (stb_image.h:2232-2330)
//   * chroma up-sampling: 3:1 / 9:3:3:1 "centred" triangle filters for 2x1, 1x2 and 2x2, nearest for anything else,
//     evaluated per output row from a near and a far source row (stb_image.h:3240-3303, :3421-3431, :3690-3705)
//   * YCbCr -> RGB in 20-bit fixed point with the Cb->G product truncated to 16 bits (stb_image.h:3434-3460)
//   * coefficients are 16-bit (dequantised products wrap to short like stb's, :2036, :2066, :2877-2882)
// The decoder below is written from the JPEG standard (ITU T.81) with those choices; it is not a copy of stb's code
// (own bit reader, canonical-code Huffman tables, block store and scan driver). tests/test_io.py checks it bit for bit
// against the targets the unmodified reference decoded (tests/golden/c1_target_rgb8.npz: progressive 4:4:4,
// c2_target_rgb8.npz: baseline 4:2:0).
// Supported: baseline and progressive Huffman, 8-bit, 1 or 3 components, sampling factors 1..4, restart intervals.
// Not supported (GS2D_ERR_FORMAT): arithmetic coding, lossless, 12-bit, 4-component (CMYK/YCCK) files.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "../../include/gs2d.h"

namespace
{
// zig-zag position -> natural (row-major) position, T.81 figure A.6
const uint8_t ZIGZAG[64] = { 0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
	                         41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
	                         30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63 };

struct Huffman
{
	// canonical code tables of T.81 annex C / F.2.2.3: codes of length L are [mincode[L], maxcode[L]]
	int maxcode[18];
	int valptr[17];
	int mincode[17];
	uint8_t values[256];
	uint8_t look_len[512]; // 9-bit lookahead: code length (0 = longer than 9 bits) and symbol
	uint8_t look_sym[512];
	bool present = false;

	bool build( const uint8_t counts[16], const uint8_t* vals, int n )
	{
		if( n > 256 )
			return false;
		memcpy( values, vals, n );
		memset( look_len, 0, sizeof( look_len ) );
		int code = 0, k = 0;
		for( int len = 1; len <= 16; len++ )
		{
			valptr[len] = k;
			mincode[len] = code;
			for( int i = 0; i < counts[len - 1]; i++, k++, code++ )
			{
				if( len <= 9 )
				{
					const int first = code << ( 9 - len );
					for( int f = 0; f < ( 1 << ( 9 - len ) ); f++ )
					{
						look_len[first + f] = (uint8_t)len;
						look_sym[first + f] = values[k];
					}
				}
			}
			maxcode[len] = counts[len - 1] ? code - 1 : -1;
			if( code > ( 1 << len ) )
				return false; // over-subscribed
			code <<= 1;
		}
		maxcode[17] = 0x7fffffff;
		present = true;
		return true;
	}
};

struct Component
{
	int id = 0, h = 1, v = 1, tq = 0;
	int hd = 0, ha = 0;       // Huffman table selectors of the current scan
	int dc_pred = 0;
	int x = 0, y = 0;         // effective size in samples
	int w2 = 0, h2 = 0;       // allocated size: whole MCUs
	int bw = 0;               // blocks per row of the allocation
	std::vector<uint8_t> data;  // w2 * h2 samples after the IDCT
	std::vector<int16_t> coeff; // progressive: 64 coefficients per block, natural order
};

struct Decoder
{
	const uint8_t* p;
	const uint8_t* end;

	// entropy-coded segment reader: MSB-first bit buffer; a marker inside the segment stops the byte supply (zeros follow)
	uint32_t bitbuf = 0;
	int bitcnt = 0;
	int marker = -1;
	bool nomore = false;

	Huffman hdc[4], hac[4];
	uint16_t dequant[4][64]; // natural order
	bool have_q[4] = { false, false, false, false };
	Component comp[3];
	int ncomp = 0;
	int W = 0, H = 0;
	bool progressive = false, have_frame = false;
	int hmax = 1, vmax = 1, mcu_x = 0, mcu_y = 0;
	int restart_interval = 0, todo = 0;
	int jfif = 0, adobe_transform = -1, rgb_ids = 0;
	// scan parameters
	int scan_n = 0, order[3] = { 0, 0, 0 };
	int ss = 0, se = 63, ah = 0, al = 0;
	int eob_run = 0;

	int get8() { return p < end ? *p++ : 0; }
	int get16()
	{
		const int a = get8();
		return ( a << 8 ) | get8();
	}

	void fill()
	{
		while( bitcnt <= 24 )
		{
			int b = nomore ? 0 : get8();
			if( b == 0xff && !nomore )
			{
				int c = get8();
				while( c == 0xff )
					c = get8();
				if( c != 0 )
				{
					marker = c;
					nomore = true;
					b = 0;
				}
			}
			bitbuf |= (uint32_t)b << ( 24 - bitcnt );
			bitcnt += 8;
		}
	}
	int bits( int n ) // n in 0..16
	{
		if( n == 0 )
			return 0;
		if( bitcnt < n )
			fill();
		const int v = (int)( bitbuf >> ( 32 - n ) );
		bitbuf <<= n;
		bitcnt -= n;
		return v;
	}
	int bit() { return bits( 1 ); }
	// RECEIVE + EXTEND, T.81 F.2.2.1 / figure F.12
	int receive_extend( int n )
	{
		const int v = bits( n );
		return v < ( 1 << ( n - 1 ) ) ? v - ( 1 << n ) + 1 : v;
	}
	int decode_symbol( const Huffman& h )
	{
		if( bitcnt < 16 )
			fill();
		const int look = (int)( bitbuf >> 23 );
		int len = h.look_len[look];
		if( len )
		{
			bitbuf <<= len;
			bitcnt -= len;
			return h.look_sym[look];
		}
		const int code16 = (int)( bitbuf >> 16 );
		for( len = 10; len <= 16; len++ )
			if( h.maxcode[len] >= 0 && ( code16 >> ( 16 - len ) ) <= h.maxcode[len] &&
			    ( code16 >> ( 16 - len ) ) >= h.mincode[len] )
				break;
		if( len > 16 )
			return -1;
		const int code = code16 >> ( 16 - len );
		bitbuf <<= len;
		bitcnt -= len;
		return h.values[h.valptr[len] + code - h.mincode[len]];
	}
	void reset_entropy()
	{
		bitbuf = 0;
		bitcnt = 0;
		nomore = false;
		marker = -1;
		for( int i = 0; i < 3; i++ )
			comp[i].dc_pred = 0;
		todo = restart_interval ? restart_interval : 0x7fffffff;
		eob_run = 0;
	}

	bool parse_tables( int m );
	bool parse_frame( int m );
	bool parse_scan_header();
	bool decode_scan();
	bool block_baseline( Component& c, int16_t d[64] );
	bool block_prog_dc( Component& c, int16_t* d );
	bool block_prog_ac( Component& c, int16_t* d );
	bool restart_check();
	void finish_progressive();
	int next_marker();
	int run();
};

// ------------------------------------------------------------------------------------------------------ IDCT ---
// One 8-point pass; inputs s0..s7, outputs e0..e3 (even part) and o0..o3 (odd part), constants scaled by 4096.
inline int fix( double x ) { return (int)( x * 4096 + 0.5 ); }
struct Pass
{
	int e0, e1, e2, e3, o0, o1, o2, o3;
};
inline Pass idct8( int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7 )
{
	static const int C0 = fix( 0.5411961f ), C1 = fix( -1.847759065f ), C2 = fix( 0.765366865f ),
	                 C3 = fix( 1.175875602f ), C4 = fix( 0.298631336f ), C5 = fix( 2.053119869f ),
	                 C6 = fix( 3.072711026f ), C7 = fix( 1.501321110f ), C8 = fix( -0.899976223f ),
	                 C9 = fix( -2.562915447f ), C10 = fix( -1.961570560f ), C11 = fix( -0.390180644f );
	Pass r;
	const int z = ( s2 + s6 ) * C0;
	const int a2 = z + s6 * C1, a3 = z + s2 * C2;
	const int a0 = ( s0 + s4 ) * 4096, a1 = ( s0 - s4 ) * 4096;
	r.e0 = a0 + a3;
	r.e3 = a0 - a3;
	r.e1 = a1 + a2;
	r.e2 = a1 - a2;
	const int q3 = s7 + s3, q4 = s5 + s1, q1 = s7 + s1, q2 = s5 + s3;
	const int z5 = ( q3 + q4 ) * C3;
	const int m1 = z5 + q1 * C8, m2 = z5 + q2 * C9, m3 = q3 * C10, m4 = q4 * C11;
	r.o3 = s1 * C7 + m1 + m4;
	r.o2 = s3 * C6 + m2 + m3;
	r.o1 = s5 * C5 + m2 + m4;
	r.o0 = s7 * C4 + m1 + m3;
	return r;
}
inline uint8_t clamp8( int x ) { return (uint8_t)( x < 0 ? 0 : ( x > 255 ? 255 : x ) ); }

void idct_block( uint8_t* out, int stride, const int16_t d[64] )
{
	int v[64];
	for( int i = 0; i < 8; i++ ) // columns, keep 2 extra bits
	{
		const Pass r = idct8( d[i], d[8 + i], d[16 + i], d[24 + i], d[32 + i], d[40 + i], d[48 + i], d[56 + i] );
		const int e0 = r.e0 + 512, e1 = r.e1 + 512, e2 = r.e2 + 512, e3 = r.e3 + 512;
		v[i] = ( e0 + r.o3 ) >> 10;
		v[56 + i] = ( e0 - r.o3 ) >> 10;
		v[8 + i] = ( e1 + r.o2 ) >> 10;
		v[48 + i] = ( e1 - r.o2 ) >> 10;
		v[16 + i] = ( e2 + r.o1 ) >> 10;
		v[40 + i] = ( e2 - r.o1 ) >> 10;
		v[24 + i] = ( e3 + r.o0 ) >> 10;
		v[32 + i] = ( e3 - r.o0 ) >> 10;
	}
	for( int i = 0; i < 8; i++ ) // rows: remove 2^17, round, level shift by 128
	{
		const int* w = v + 8 * i;
		const Pass r = idct8( w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7] );
		const int bias = 65536 + ( 128 << 17 );
		const int e0 = r.e0 + bias, e1 = r.e1 + bias, e2 = r.e2 + bias, e3 = r.e3 + bias;
		uint8_t* o = out + (size_t)stride * i;
		o[0] = clamp8( ( e0 + r.o3 ) >> 17 );
		o[7] = clamp8( ( e0 - r.o3 ) >> 17 );
		o[1] = clamp8( ( e1 + r.o2 ) >> 17 );
		o[6] = clamp8( ( e1 - r.o2 ) >> 17 );
		o[2] = clamp8( ( e2 + r.o1 ) >> 17 );
		o[5] = clamp8( ( e2 - r.o1 ) >> 17 );
		o[3] = clamp8( ( e3 + r.o0 ) >> 17 );
		o[4] = clamp8( ( e3 - r.o0 ) >> 17 );
	}
}
} // namespace

namespace
{
// ----------------------------------------------------------------------------------------- marker segments ---
// DQT, DHT, DRI, APPn, COM. Returns false on a malformed or unsupported segment.
bool Decoder::parse_tables( int m )
{
	if( m == 0xDD ) // DRI
	{
		if( get16() != 4 )
			return false;
		restart_interval = get16();
		return true;
	}
	if( m == 0xDB ) // DQT
	{
		int L = get16() - 2;
		while( L > 0 )
		{
			const int q = get8();
			const int prec16 = q >> 4, t = q & 15;
			if( prec16 > 1 || t > 3 )
				return false;
			for( int i = 0; i < 64; i++ )
				dequant[t][ZIGZAG[i]] = (uint16_t)( prec16 ? get16() : get8() );
			have_q[t] = true;
			L -= prec16 ? 129 : 65;
		}
		return L == 0;
	}
	if( m == 0xC4 ) // DHT
	{
		int L = get16() - 2;
		while( L > 0 )
		{
			const int q = get8();
			const int tc = q >> 4, th = q & 15;
			if( tc > 1 || th > 3 )
				return false;
			uint8_t counts[16], vals[256];
			int n = 0;
			for( int i = 0; i < 16; i++ )
			{
				counts[i] = (uint8_t)get8();
				n += counts[i];
			}
			if( n > 256 )
				return false;
			L -= 17;
			for( int i = 0; i < n; i++ )
				vals[i] = (uint8_t)get8();
			L -= n;
			if( !( tc == 0 ? hdc[th] : hac[th] ).build( counts, vals, n ) )
				return false;
		}
		return L == 0;
	}
	if( ( m >= 0xE0 && m <= 0xEF ) || m == 0xFE ) // APPn, COM
	{
		int L = get16();
		if( L < 2 )
			return false;
		L -= 2;
		if( m == 0xE0 && L >= 5 )
		{
			static const char tag[5] = { 'J', 'F', 'I', 'F', 0 };
			bool ok = true;
			for( int i = 0; i < 5; i++ )
				ok &= get8() == tag[i];
			L -= 5;
			if( ok )
				jfif = 1;
		}
		else if( m == 0xEE && L >= 12 )
		{
			static const char tag[6] = { 'A', 'd', 'o', 'b', 'e', 0 };
			bool ok = true;
			for( int i = 0; i < 6; i++ )
				ok &= get8() == tag[i];
			L -= 6;
			if( ok )
			{
				get8();  // version
				get16(); // flags0
				get16(); // flags1
				adobe_transform = get8();
				L -= 6;
			}
		}
		if( L > end - p )
			return false;
		p += L;
		return true;
	}
	return false;
}

bool Decoder::parse_frame( int m )
{
	if( m != 0xC0 && m != 0xC1 && m != 0xC2 )
		return false; // lossless / arithmetic / hierarchical
	progressive = m == 0xC2;
	const int Lf = get16();
	if( Lf < 11 || get8() != 8 )
		return false;
	H = get16();
	W = get16();
	const int c = get8();
	if( H == 0 || W == 0 || ( c != 1 && c != 3 ) || Lf != 8 + 3 * c )
		return false;
	ncomp = c;
	rgb_ids = 0;
	static const char rgb[3] = { 'R', 'G', 'B' };
	for( int i = 0; i < c; i++ )
	{
		comp[i].id = get8();
		if( c == 3 && comp[i].id == rgb[i] )
			rgb_ids++;
		const int q = get8();
		comp[i].h = q >> 4;
		comp[i].v = q & 15;
		comp[i].tq = get8();
		if( comp[i].h < 1 || comp[i].h > 4 || comp[i].v < 1 || comp[i].v > 4 || comp[i].tq > 3 )
			return false;
	}
	if( (uint64_t)W * H > ( 1ull << 30 ) )
		return false;
	hmax = vmax = 1;
	for( int i = 0; i < c; i++ )
	{
		hmax = comp[i].h > hmax ? comp[i].h : hmax;
		vmax = comp[i].v > vmax ? comp[i].v : vmax;
	}
	mcu_x = ( W + hmax * 8 - 1 ) / ( hmax * 8 );
	mcu_y = ( H + vmax * 8 - 1 ) / ( vmax * 8 );
	for( int i = 0; i < c; i++ )
	{
		Component& k = comp[i];
		k.x = ( W * k.h + hmax - 1 ) / hmax;
		k.y = ( H * k.v + vmax - 1 ) / vmax;
		k.w2 = mcu_x * k.h * 8;
		k.h2 = mcu_y * k.v * 8;
		k.bw = k.w2 / 8;
		k.data.assign( (size_t)k.w2 * k.h2, 0 );
		if( progressive )
			k.coeff.assign( (size_t)k.w2 * k.h2, 0 );
	}
	have_frame = true;
	return true;
}

bool Decoder::parse_scan_header()
{
	const int Ls = get16();
	scan_n = get8();
	if( scan_n < 1 || scan_n > ncomp || Ls != 6 + 2 * scan_n )
		return false;
	for( int i = 0; i < scan_n; i++ )
	{
		const int id = get8(), q = get8();
		int which = 0;
		while( which < ncomp && comp[which].id != id )
			which++;
		if( which == ncomp )
			return false;
		comp[which].hd = q >> 4;
		comp[which].ha = q & 15;
		if( comp[which].hd > 3 || comp[which].ha > 3 )
			return false;
		order[i] = which;
	}
	ss = get8();
	se = get8();
	const int a = get8();
	ah = a >> 4;
	al = a & 15;
	if( progressive )
	{
		if( ss > 63 || se > 63 || ss > se || ah > 13 || al > 13 )
			return false;
	}
	else
	{
		if( ss != 0 || ah != 0 || al != 0 )
			return false;
		se = 63;
	}
	return true;
}
} // namespace

namespace
{
// ------------------------------------------------------------------------------------------- block decoders ---
// Sequential block (T.81 F.2.2): DC difference + run/size coded AC; dequantised while decoding, 16-bit products.
bool Decoder::block_baseline( Component& c, int16_t d[64] )
{
	const Huffman &dc = hdc[c.hd], &ac = hac[c.ha];
	const uint16_t* q = dequant[c.tq];
	const int t = decode_symbol( dc );
	if( t < 0 || t > 16 )
		return false;
	memset( d, 0, 64 * sizeof( int16_t ) );
	c.dc_pred += t ? receive_extend( t ) : 0;
	d[0] = (int16_t)( c.dc_pred * q[0] );
	int k = 1;
	while( k < 64 )
	{
		const int rs = decode_symbol( ac );
		if( rs < 0 )
			return false;
		const int s = rs & 15, r = rs >> 4;
		if( s == 0 )
		{
			if( rs != 0xf0 )
				break; // EOB
			k += 16;
			continue;
		}
		k += r;
		if( k > 63 )
			return false;
		const int nat = ZIGZAG[k++];
		d[nat] = (int16_t)( receive_extend( s ) * q[nat] );
	}
	return true;
}

// Progressive DC scan (T.81 G.1.2.1): first pass stores the point-transformed DC, refinement passes add one bit.
bool Decoder::block_prog_dc( Component& c, int16_t* d )
{
	if( se != 0 )
		return false;
	if( ah == 0 )
	{
		memset( d, 0, 64 * sizeof( int16_t ) );
		const int t = decode_symbol( hdc[c.hd] );
		if( t < 0 || t > 16 )
			return false;
		c.dc_pred += t ? receive_extend( t ) : 0;
		d[0] = (int16_t)( c.dc_pred << al );
	}
	else if( bit() )
		d[0] = (int16_t)( d[0] + ( 1 << al ) );
	return true;
}

// Progressive AC scan (T.81 G.1.2.2 first pass with EOB runs, G.1.2.3 refinement with correction bits).
bool Decoder::block_prog_ac( Component& c, int16_t* d )
{
	if( ss == 0 )
		return false;
	const Huffman& ac = hac[c.ha];
	if( ah == 0 )
	{
		if( eob_run )
		{
			eob_run--;
			return true;
		}
		int k = ss;
		while( k <= se )
		{
			const int rs = decode_symbol( ac );
			if( rs < 0 )
				return false;
			const int s = rs & 15, r = rs >> 4;
			if( s == 0 )
			{
				if( r < 15 )
				{
					eob_run = ( 1 << r ) - 1;
					if( r )
						eob_run += bits( r );
					break;
				}
				k += 16;
				continue;
			}
			k += r;
			if( k > 63 )
				return false;
			d[ZIGZAG[k++]] = (int16_t)( receive_extend( s ) << al );
		}
		return true;
	}
	// refinement
	const int16_t one = (int16_t)( 1 << al );
	auto correct = [&]( int16_t& v ) { // a correction bit for an already non-zero coefficient
		if( bit() && ( v & one ) == 0 )
			v = (int16_t)( v > 0 ? v + one : v - one );
	};
	if( eob_run )
	{
		eob_run--;
		for( int k = ss; k <= se; k++ )
		{
			int16_t& v = d[ZIGZAG[k]];
			if( v != 0 )
				correct( v );
		}
		return true;
	}
	int k = ss;
	while( k <= se )
	{
		const int rs = decode_symbol( ac );
		if( rs < 0 )
			return false;
		int s = rs & 15, r = rs >> 4;
		int newval = 0;
		if( s == 0 )
		{
			if( r < 15 )
			{
				eob_run = ( 1 << r ) - 1;
				if( r )
					eob_run += bits( r );
				r = 64; // run to the end of the band, correcting the non-zero coefficients on the way
			}
		}
		else
		{
			if( s != 1 )
				return false;
			newval = bit() ? one : -one;
		}
		while( k <= se )
		{
			int16_t& v = d[ZIGZAG[k++]];
			if( v != 0 )
				correct( v );
			else
			{
				if( r == 0 )
				{
					v = (int16_t)newval;
					break;
				}
				r--;
			}
		}
	}
	return true;
}

// after every MCU: count down the restart interval; at 0 expect RSTn, else the scan ends here
bool Decoder::restart_check()
{
	if( --todo > 0 )
		return true;
	if( bitcnt < 24 )
		fill();
	if( marker < 0xD0 || marker > 0xD7 )
		return false;
	reset_entropy();
	return true;
}

bool Decoder::decode_scan()
{
	reset_entropy();
	for( int i = 0; i < scan_n; i++ )
	{
		const Component& c = comp[order[i]];
		const bool need_dc = !progressive || ss == 0, need_ac = !progressive || ss > 0;
		if( ( need_dc && ah == 0 && !hdc[c.hd].present ) || ( need_ac && !hac[c.ha].present ) )
			return false;
		if( !have_q[c.tq] )
			return false;
	}
	int16_t blk[64];
	if( scan_n == 1 )
	{
		// non-interleaved: the component's own blocks in raster order, every block is an MCU
		Component& c = comp[order[0]];
		const int bw = ( c.x + 7 ) >> 3, bh = ( c.y + 7 ) >> 3;
		for( int by = 0; by < bh; by++ )
			for( int bx = 0; bx < bw; bx++ )
			{
				if( !progressive )
				{
					if( !block_baseline( c, blk ) )
						return false;
					idct_block( c.data.data() + (size_t)c.w2 * by * 8 + bx * 8, c.w2, blk );
				}
				else
				{
					int16_t* d = c.coeff.data() + 64 * ( (size_t)bx + (size_t)by * c.bw );
					if( !( ss == 0 ? block_prog_dc( c, d ) : block_prog_ac( c, d ) ) )
						return false;
				}
				if( !restart_check() )
					return true;
			}
		return true;
	}
	// interleaved MCUs: h x v blocks of every scan component
	for( int my = 0; my < mcu_y; my++ )
		for( int mx = 0; mx < mcu_x; mx++ )
		{
			for( int i = 0; i < scan_n; i++ )
			{
				Component& c = comp[order[i]];
				for( int y = 0; y < c.v; y++ )
					for( int x = 0; x < c.h; x++ )
					{
						const int bx = mx * c.h + x, by = my * c.v + y;
						if( !progressive )
						{
							if( !block_baseline( c, blk ) )
								return false;
							idct_block( c.data.data() + (size_t)c.w2 * by * 8 + bx * 8, c.w2, blk );
						}
						else
						{
							// interleaved progressive scans carry DC only (T.81 G.1.1.1.1)
							if( !block_prog_dc( c, c.coeff.data() + 64 * ( (size_t)bx + (size_t)by * c.bw ) ) )
								return false;
						}
					}
			}
			if( !restart_check() )
				return true;
		}
	return true;
}

void Decoder::finish_progressive()
{
	for( int n = 0; n < ncomp; n++ )
	{
		Component& c = comp[n];
		const uint16_t* q = dequant[c.tq];
		const int bw = ( c.x + 7 ) >> 3, bh = ( c.y + 7 ) >> 3;
		for( int by = 0; by < bh; by++ )
			for( int bx = 0; bx < bw; bx++ )
			{
				int16_t* d = c.coeff.data() + 64 * ( (size_t)bx + (size_t)by * c.bw );
				for( int i = 0; i < 64; i++ )
					d[i] = (int16_t)( d[i] * q[i] );
				idct_block( c.data.data() + (size_t)c.w2 * by * 8 + bx * 8, c.w2, d );
			}
	}
}

// the marker the entropy reader ran into, else the next one in the stream (0xFF fill bytes skipped); -1 = none
int Decoder::next_marker()
{
	if( marker >= 0 )
	{
		const int m = marker;
		marker = -1;
		return m;
	}
	int x = get8();
	if( x != 0xff )
		return -1;
	while( x == 0xff )
		x = get8();
	return x;
}
} // namespace

namespace
{
// -------------------------------------------------------------------------------- up-sampling and colour ---
// One output row of a component from its `near` and `far` source rows (the vertical phase decides which is which).
// hs x vs = 2x2: 9:3:3:1 filter; 2x1: 3:1 horizontally; 1x2: 3:1 vertically; 1x1: the row itself; else nearest.
const uint8_t* upsample_row( uint8_t* out, const uint8_t* near, const uint8_t* far, int w, int hs, int vs )
{
	if( hs == 1 && vs == 1 )
		return near;
	if( hs == 1 && vs == 2 )
	{
		for( int i = 0; i < w; i++ )
			out[i] = (uint8_t)( ( 3 * near[i] + far[i] + 2 ) >> 2 );
		return out;
	}
	if( hs == 2 && vs == 1 )
	{
		if( w == 1 )
		{
			out[0] = out[1] = near[0];
			return out;
		}
		out[0] = near[0];
		out[1] = (uint8_t)( ( near[0] * 3 + near[1] + 2 ) >> 2 );
		for( int i = 1; i < w - 1; i++ )
		{
			const int n = 3 * near[i] + 2;
			out[2 * i] = (uint8_t)( ( n + near[i - 1] ) >> 2 );
			out[2 * i + 1] = (uint8_t)( ( n + near[i + 1] ) >> 2 );
		}
		out[2 * ( w - 1 )] = (uint8_t)( ( near[w - 2] * 3 + near[w - 1] + 2 ) >> 2 );
		out[2 * ( w - 1 ) + 1] = near[w - 1];
		return out;
	}
	if( hs == 2 && vs == 2 )
	{
		// vertical 3:1 first (kept at 4x scale), then horizontal 3:1 -> one rounding by 16
		int cur = 3 * near[0] + far[0];
		if( w == 1 )
		{
			out[0] = out[1] = (uint8_t)( ( cur + 2 ) >> 2 );
			return out;
		}
		out[0] = (uint8_t)( ( cur + 2 ) >> 2 );
		for( int i = 1; i < w; i++ )
		{
			const int prev = cur;
			cur = 3 * near[i] + far[i];
			out[2 * i - 1] = (uint8_t)( ( 3 * prev + cur + 8 ) >> 4 );
			out[2 * i] = (uint8_t)( ( 3 * cur + prev + 8 ) >> 4 );
		}
		out[2 * w - 1] = (uint8_t)( ( cur + 2 ) >> 2 );
		return out;
	}
	for( int i = 0; i < w; i++ )
		for( int j = 0; j < hs; j++ )
			out[i * hs + j] = near[i];
	return out;
}

inline int fixed20( float x ) { return ( (int)( x * 4096.0f + 0.5f ) ) << 8; }

void ycc_to_rgba_row( uint8_t* out, const uint8_t* y, const uint8_t* pcb, const uint8_t* pcr, int n )
{
	static const int KR = fixed20( 1.40200f ), KGR = fixed20( 0.71414f ), KGB = fixed20( 0.34414f ),
	                 KB = fixed20( 1.77200f );
	for( int i = 0; i < n; i++ )
	{
		const int yf = ( y[i] << 20 ) + ( 1 << 19 ); // + 0.5 for the final truncation
		const int cr = pcr[i] - 128, cb = pcb[i] - 128;
		int r = yf + cr * KR;
		// the Cb contribution to green is truncated to a multiple of 2^16 before it is added (stb_image.h:3444)
		int g = (int)( (uint32_t)( yf + cr * -KGR ) + ( (uint32_t)( cb * -KGB ) & 0xffff0000u ) );
		int b = yf + cb * KB;
		r >>= 20;
		g >>= 20;
		b >>= 20;
		out[4 * i + 0] = clamp8( r );
		out[4 * i + 1] = clamp8( g );
		out[4 * i + 2] = clamp8( b );
		out[4 * i + 3] = 255;
	}
}

// 0 = ok, else GS2D_ERR_FORMAT
int Decoder::run()
{
	if( get8() != 0xff || get8() != 0xd8 ) // SOI
		return GS2D_ERR_FORMAT;
	bool seen_scan = false;
	for( ;; )
	{
		int m = next_marker();
		if( m == 0xD9 ) // EOI
			break;
		if( m == -1 )
		{
			// junk or zero padding between segments: resynchronise on the next 0xFF <non-zero> (stb_image.h:3200-3212)
			bool found = false;
			while( p < end )
			{
				if( get8() == 0xff )
				{
					const int x = get8();
					if( x != 0 && x != 0xff )
					{
						marker = x;
						found = true;
						break;
					}
				}
			}
			if( !found )
				break; // data ended without EOI: what has been decoded is the picture
			continue;
		}
		if( m == 0xC0 || m == 0xC1 || m == 0xC2 )
		{
			if( have_frame || !parse_frame( m ) )
				return GS2D_ERR_FORMAT;
		}
		else if( ( m >= 0xC3 && m <= 0xCF && m != 0xC4 ) )
			return GS2D_ERR_FORMAT; // lossless, hierarchical or arithmetic-coded frame
		else if( m == 0xDA )
		{
			if( !have_frame || !parse_scan_header() || !decode_scan() )
				return GS2D_ERR_FORMAT;
			seen_scan = true;
		}
		else if( m == 0xDC ) // DNL
		{
			const int Ld = get16();
			const int NL = get16();
			if( Ld != 4 || NL != H )
				return GS2D_ERR_FORMAT;
		}
		else if( !parse_tables( m ) )
			return GS2D_ERR_FORMAT;
	}
	if( !have_frame || !seen_scan )
		return GS2D_ERR_FORMAT;
	if( progressive )
		finish_progressive();
	return GS2D_OK;
}
} // namespace

extern "C" int gs2d_image_decode_jpeg_rgba8( const uint8_t* bytes, size_t n_bytes, uint8_t** rgba, int* W, int* H )
{
	if( !bytes || !rgba || !W || !H || n_bytes < 4 )
		return GS2D_ERR_INVALID;
	Decoder* dec = new Decoder(); // ~8 KB of tables: not on the stack
	dec->p = bytes;
	dec->end = bytes + n_bytes;
	int rc = dec->run();
	if( rc != GS2D_OK )
	{
		delete dec;
		return rc;
	}
	const int w = dec->W, h = dec->H;
	uint8_t* out = (uint8_t*)malloc( (size_t)w * h * 4 );
	if( !out )
	{
		delete dec;
		return GS2D_ERR_OOM;
	}
	// R'G'B' stored directly: component ids "RGB", or an Adobe segment with transform 0 and no JFIF header
	const bool is_rgb = dec->ncomp == 3 && ( dec->rgb_ids == 3 || ( dec->adobe_transform == 0 && !dec->jfif ) );
	struct Up
	{
		int hs, vs, ystep, ypos, w_lores;
		const uint8_t *line0, *line1;
		std::vector<uint8_t> buf;
	} up[3];
	for( int k = 0; k < dec->ncomp; k++ )
	{
		Component& c = dec->comp[k];
		up[k].hs = dec->hmax / c.h;
		up[k].vs = dec->vmax / c.v;
		up[k].ystep = up[k].vs >> 1;
		up[k].ypos = 0;
		up[k].w_lores = ( w + up[k].hs - 1 ) / up[k].hs;
		up[k].line0 = up[k].line1 = c.data.data();
		up[k].buf.resize( (size_t)w + 8 );
	}
	for( int j = 0; j < h; j++ )
	{
		const uint8_t* row[3] = { nullptr, nullptr, nullptr };
		for( int k = 0; k < dec->ncomp; k++ )
		{
			Up& u = up[k];
			// the output row sits in the lower half of its source row <=> the nearer neighbour is the row below
			const bool below = u.ystep >= ( u.vs >> 1 );
			row[k] = upsample_row( u.buf.data(), below ? u.line1 : u.line0, below ? u.line0 : u.line1, u.w_lores, u.hs,
			                       u.vs );
			if( ++u.ystep >= u.vs )
			{
				u.ystep = 0;
				u.line0 = u.line1;
				if( ++u.ypos < dec->comp[k].y )
					u.line1 += dec->comp[k].w2;
			}
		}
		uint8_t* o = out + (size_t)j * w * 4;
		if( dec->ncomp == 1 )
			for( int i = 0; i < w; i++ )
			{
				o[4 * i] = o[4 * i + 1] = o[4 * i + 2] = row[0][i];
				o[4 * i + 3] = 255;
			}
		else if( is_rgb )
			for( int i = 0; i < w; i++ )
			{
				o[4 * i] = row[0][i];
				o[4 * i + 1] = row[1][i];
				o[4 * i + 2] = row[2][i];
				o[4 * i + 3] = 255;
			}
		else
			ycc_to_rgba_row( o, row[0], row[1], row[2], w );
	}
	delete dec;
	*rgba = out;
	*W = w;
	*H = h;
	return GS2D_OK;
}
