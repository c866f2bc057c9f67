/* oracle/gs2d_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE (see gs2d_oracle.h).
 *
 * Plain-C restatement of /root/reference/main.cpp:414-809. Every expression keeps the reference's operand order
 * and is evaluated in float without contraction (compile with -ffp-contract=off). glm facts used, all from the
 * reference's vendored glm 0.9.9.6 under libs/prlib/src/glm/detail:
 *   mat2(x0,y0,x1,y1) has columns (x0,y0),(x1,y1), m[c][r]            type_mat2x2.inl:49-53
 *   M*v = (m00*vx + m10*vy, m01*vx + m11*vy)                           type_mat2x2.inl:429-438
 *   M/s divides every element by s                                     type_mat2x2.inl:489-494
 *   dot(a,b) = sum of the component products, left to right            func_geometric.inl:38-57
 *   mix(x,y,a) = x*(1-a) + y*a                                         func_common.inl:123-133
 *   clamp(x,lo,hi) = min(max(x,lo),hi)                                 func_common.inl:504-508
 */
#include "gs2d_oracle.h"

#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* float -> int as the reference's x86 build performs it (cvttss2si): truncation toward zero, and the "integer
 * indefinite" INT_MIN for NaN / out-of-range (the C cast is undefined there; main.cpp:490-491, :507-508). */
static int trunc_x86( float f )
{
	if( !( f > -2147483648.0f && f < 2147483648.0f ) )
		return INT_MIN;
	return (int)f;
}

static float sign_of( float v ) { return v < 0.0f ? -1.0f : 1.0f; } /* main.cpp:34-37 */
static float ss_max( float x, float y ) { return ( x < y ) ? y : x; } /* main.cpp:38-42 */
static float ss_min( float x, float y ) { return ( y < x ) ? y : x; } /* main.cpp:44-48 */
/* glm::clamp = min( max( x, lo ), hi ) with glm's comparison-based min/max (func_common.inl:17-30, :504-508):
 * a NaN x survives both (so the finite check at main.cpp:752-785 can still see it), unlike fminf/fmaxf. */
static float clampf( float x, float lo, float hi )
{
	float t = ( x < lo ) ? lo : x;
	return ( hi < t ) ? hi : t;
}
static float mixf( float x, float y, float a ) { return x * ( 1.0f - a ) + y * a; }

/* main.cpp:49-83 */
static float exp_approx( float x )
{
	x = 1.0f + x / 8.0f;
	if( x < 0.00001814586175896693021059036255f )
		return 0.0f;
	x *= x;
	x *= x;
	x *= x;
	return x;
}

/* main.cpp:171-185 */
static int solve_quadratic( float xs[2], float a, float b, float c )
{
	float det = b * b - 4.0f * a * c;
	if( det < 0.0f )
		return 0;
	float k = ( -b - sign_of( b ) * sqrtf( det ) ) / 2.0f;
	float x0 = k / a;
	float x1 = c / k;
	xs[0] = ss_min( x0, x1 );
	xs[1] = ss_max( x0, x1 );
	return 2;
}

void oracle_pcg3d( uint32_t v[3] )
{
	v[0] = v[0] * 1664525u + 1013904223u;
	v[1] = v[1] * 1664525u + 1013904223u;
	v[2] = v[2] * 1664525u + 1013904223u;
	v[0] += v[1] * v[2];
	v[1] += v[2] * v[0];
	v[2] += v[0] * v[1];
	v[0] ^= v[0] >> 16u;
	v[1] ^= v[1] >> 16u;
	v[2] ^= v[2] >> 16u;
	v[0] += v[1] * v[2];
	v[1] += v[2] * v[0];
	v[2] += v[0] * v[1];
}

void oracle_init_splats( int W, int H, int n, gs2d_splat* splats )
{
	/* glm::vec3( 0xFFFFFFFF, ... ): the unsigned literal converts to float 2^32 (main.cpp:290-291) */
	const float denom = (float)0xFFFFFFFFu;
	for( int i = 0; i < n; i++ )
	{
		uint32_t h0[3] = { (uint32_t)i, 0u, 0xFFFFFFFFu };
		uint32_t h1[3] = { (uint32_t)i, 1u, 0xFFFFFFFFu };
		oracle_pcg3d( h0 );
		oracle_pcg3d( h1 );
		float r0x = (float)h0[0] / denom, r0y = (float)h0[1] / denom;
		float r1x = (float)h1[0] / denom, r1y = (float)h1[1] / denom, r1z = (float)h1[2] / denom;
		gs2d_splat s;
		s.pos[0] = mixf( r0x, (float)W - 1, r0x ); /* sic: mix( r0.x, w-1, r0.x ), main.cpp:294 */
		s.pos[1] = mixf( r0y, (float)H - 1, r0y );
		s.sx = mixf( 6.0f, 10.0f, r1x );
		s.sy = mixf( 6.0f, 10.0f, r1y );
		s.rot = 3.14159265358979323846264338327950288f * r1z;
		s.color[0] = s.color[1] = s.color[2] = 0.5f;
		s.opacity = 1.0f;
		splats[i] = s;
	}
}

/* Per-splat quantities both passes derive first (cov_of main.cpp:206-221; inverse :432-436 / :560-565). */
typedef struct
{
	float a, b, c, d; /* inv_cov[0][0], inv_cov[1][0], inv_cov[0][1], inv_cov[1][1] */
	float det;
	float cosT, sinT;
} splat_geom;

static splat_geom geom_of( const gs2d_splat* s )
{
	splat_geom g;
	float theta = s->rot;
	float sx = s->sx, sy = s->sy;
	float cosTheta = cosf( theta );
	float sinTheta = sinf( theta );
	float lambda0 = sx * sx;
	float lambda1 = sy * sy;
	float s11 = lambda0 * cosTheta * cosTheta + lambda1 * sinTheta * sinTheta;
	float s12 = ( lambda0 - lambda1 ) * sinTheta * cosTheta;
	float s22 = lambda0 + lambda1 - s11;
	/* cov = mat2( s11, s12, s12, s22 ): cov[0][0]=s11 cov[0][1]=s12 cov[1][0]=s12 cov[1][1]=s22 */
	float det = s11 * s22 - s12 * s12; /* main.cpp:191 and :560 (same expression) */
	/* inv_cov = mat2( cov[1][1], -cov[0][1], -cov[1][0], cov[0][0] ) / det */
	g.a = s22 / det;    /* [0][0] */
	g.c = -s12 / det;   /* [0][1] */
	g.b = -s12 / det;   /* [1][0] */
	g.d = s11 / det;    /* [1][1] */
	g.det = det;
	g.cosT = cosTheta;
	g.sinT = sinTheta;
	return g;
}

/* Row range (main.cpp:489-491) and the x interval of one row (main.cpp:498-509), clipped to the image the way the
 * `continue` statements at :494 and :513 do. Returns 0 for an empty row. */
static void row_range( const gs2d_splat* s, const splat_geom* g, float bounds, int H, int* y0, int* y1 )
{
	float hsize_invCovY = sqrtf( g->a * g->det ) * bounds;
	int begY = trunc_x86( s->pos[1] - hsize_invCovY );
	int endY = trunc_x86( s->pos[1] + hsize_invCovY );
	*y0 = begY < 0 ? 0 : begY;
	*y1 = endY > H - 1 ? H - 1 : endY;
}

static int x_range( const gs2d_splat* s, const splat_geom* g, float bounds, int W, int y, int* x0, int* x1 )
{
	float vy = ( (float)y + 0.5f ) - s->pos[1];
	float xs[2];
	int begX = -1;
	int endX = -1;
	if( solve_quadratic( xs, g->a, 2.0f * g->b * vy, g->d * vy * vy - bounds * bounds ) )
	{
		begX = trunc_x86( s->pos[0] + xs[0] );
		endX = trunc_x86( s->pos[0] + xs[1] );
	}
	*x0 = begX < 0 ? 0 : begX;
	*x1 = endX > W - 1 ? W - 1 : endX;
	return *x0 <= *x1;
}

void oracle_forward( int W, int H, int n, const gs2d_splat* splats, const gs2d_params* p, float* image,
                     uint64_t counts[2] )
{
	uint64_t visited = 0, live = 0;
	for( size_t i = 0; i < (size_t)W * H; i++ ) /* main.cpp:414 */
	{
		image[4 * i + 0] = 0.0f;
		image[4 * i + 1] = 0.0f;
		image[4 * i + 2] = 0.0f;
		image[4 * i + 3] = 1.0f;
	}
	for( int i = 0; i < n; i++ )
	{
		gs2d_splat s = splats[i];
		splat_geom g = geom_of( &s );
		int y0, y1;
		row_range( &s, &g, p->bounds, H, &y0, &y1 );
		for( int y = y0; y <= y1; y++ )
		{
			int x0, x1;
			if( !x_range( &s, &g, p->bounds, W, y, &x0, &x1 ) )
				continue;
			for( int x = x0; x <= x1; x++ )
			{
				float* color = image + 4 * ( (size_t)y * W + x );
				float T = color[3];
				visited++;
				if( T < p->min_throughput )
					continue;
				live++;
				float vx = ( (float)x + 0.5f ) - s.pos[0];
				float vy = ( (float)y + 0.5f ) - s.pos[1];
				float rx = g.a * vx + g.b * vy; /* (inv_cov * v).x = m00*vx + m10*vy */
				float ry = g.c * vx + g.d * vy; /* (inv_cov * v).y = m01*vx + m11*vy */
				float d2 = vx * rx + vy * ry;
				float alpha = exp_approx( -0.5f * d2 ) * s.opacity;
				color[0] += T * s.color[0] * alpha;
				color[1] += T * s.color[1] * alpha;
				color[2] += T * s.color[2] * alpha;
				color[3] *= ( 1.0f - alpha );
			}
		}
	}
	for( size_t i = 0; i < (size_t)W * H; i++ ) /* main.cpp:543-546 */
		image[4 * i + 3] = 1.0f;
	if( counts )
	{
		counts[0] = visited;
		counts[1] = live;
	}
}

void oracle_backward( int W, int H, int n, const gs2d_splat* splats, const gs2d_params* p, const float* image,
                      const float* target, float* scratch, gs2d_splat* dsplats )
{
	for( size_t i = 0; i < (size_t)W * H; i++ ) /* main.cpp:549 */
	{
		scratch[4 * i + 0] = 0.0f;
		scratch[4 * i + 1] = 0.0f;
		scratch[4 * i + 2] = 0.0f;
		scratch[4 * i + 3] = 1.0f;
	}
	memset( dsplats, 0, sizeof( gs2d_splat ) * (size_t)n ); /* main.cpp:550 */

	for( int i = 0; i < n; i++ )
	{
		gs2d_splat s = splats[i];
		splat_geom g = geom_of( &s );
		float cosTheta = g.cosT, sinTheta = g.sinT;
		gs2d_splat* ds = &dsplats[i];
		int y0, y1;
		row_range( &s, &g, p->bounds, H, &y0, &y1 );
		for( int y = y0; y <= y1; y++ )
		{
			int x0, x1;
			if( !x_range( &s, &g, p->bounds, W, y, &x0, &x1 ) )
				continue;
			for( int x = x0; x <= x1; x++ )
			{
				size_t px = (size_t)y * W + x;
				float* color = scratch + 4 * px;
				float T = color[3];
				if( T < p->min_throughput )
					continue;
				float vx = ( (float)x + 0.5f ) - s.pos[0];
				float vy = ( (float)y + 0.5f ) - s.pos[1];
				float rx = g.a * vx + g.b * vy;
				float ry = g.c * vx + g.d * vy;
				float d2 = vx * rx + vy * ry;
				float G = exp_approx( -0.5f * d2 );
				float alpha = G * s.opacity;

				const float* finalColor = image + 4 * px;
				const float* ref = target + 4 * px;
				float dL_dC[3] = { finalColor[0] - ref[0], finalColor[1] - ref[1], finalColor[2] - ref[2] };
				{
					float dC_dc = alpha * T;
					ds->color[0] += dL_dC[0] * dC_dc;
					ds->color[1] += dL_dC[1] * dC_dc;
					ds->color[2] += dL_dC[2] * dC_dc;
				}
				color[0] += T * s.color[0] * alpha;
				color[1] += T * s.color[1] * alpha;
				color[2] += T * s.color[2] * alpha;

				float S[3] = { finalColor[0] - color[0], finalColor[1] - color[1], finalColor[2] - color[2] };
				float den = 1.0f - alpha + 1.0e-15f;
				float dC_dalpha[3] = { s.color[0] * T - S[0] / den, s.color[1] * T - S[1] / den,
				                       s.color[2] * T - S[2] / den };
				float dL_dalpha[3] = { dL_dC[0] * dC_dalpha[0], dL_dC[1] * dC_dalpha[1], dL_dC[2] * dC_dalpha[2] };
				float dL_dalpha_rgb = dL_dalpha[0] + dL_dalpha[1] + dL_dalpha[2];
				{
					float a = g.a, b = g.b, c = g.c, d = g.d;
					float dalpha_dx = 0.5f * alpha * ( 2.0f * a * vx + ( b + c ) * vy );
					float dalpha_dy = 0.5f * alpha * ( 2.0f * d * vy + ( b + c ) * vx );
					ds->pos[0] += dL_dalpha_rgb * dalpha_dx;
					ds->pos[1] += dL_dalpha_rgb * dalpha_dy;

					float q0 = vx * vx, q1 = vx * vy, q2 = vy * vy;
					float dalpha_dsx =
					    alpha / ( s.sx * s.sx * s.sx ) *
					    ( ( cosTheta * cosTheta ) * q0 + ( 2.0f * sinTheta * cosTheta ) * q1 + ( sinTheta * sinTheta ) * q2 );
					float dalpha_dsy =
					    alpha / ( s.sy * s.sy * s.sy ) *
					    ( ( sinTheta * sinTheta ) * q0 + ( -2.0f * sinTheta * cosTheta ) * q1 + ( cosTheta * cosTheta ) * q2 );
					ds->sx += dL_dalpha_rgb * dalpha_dsx;
					ds->sy += dL_dalpha_rgb * dalpha_dsy;

					float dalpha_dtheta = alpha * ( s.sx * s.sx - s.sy * s.sy ) / ( s.sx * s.sx * s.sy * s.sy ) *
					                      ( ( cosTheta * cosTheta - sinTheta * sinTheta ) * vx * vy -
					                        sinTheta * cosTheta * ( vx * vx - vy * vy ) );
					ds->rot += ( dL_dalpha[0] + dL_dalpha[1] + dL_dalpha[2] ) * dalpha_dtheta;

					float dalpha_do = G;
					ds->opacity += dL_dalpha_rgb * dalpha_do;
				}
				color[3] *= ( 1.0f - alpha );
			}
		}
	}
}

/* Adam::optimize, main.cpp:144-156 */
static float adam_optimize( gs2d_moment* st, float value, float g, float alpha, float beta1t, float beta2t,
                            const gs2d_params* p )
{
	float s = alpha;
	float m = p->beta1 * st->m + ( 1.0f - p->beta1 ) * g;
	float v = p->beta2 * st->v + ( 1.0f - p->beta2 ) * g * g;
	st->m = m;
	st->v = v;
	float m_hat = m / ( 1.0f - beta1t );
	float v_hat = v / ( 1.0f - beta2t );
	return value - s * m_hat / ( sqrtf( v_hat ) + p->eps );
}

int oracle_adam( int W, int H, int n, gs2d_splat* splats, gs2d_splat_adam* adam, const gs2d_splat* dsplats,
                 float* beta1t, float* beta2t, const gs2d_params* p )
{
	float trainingRate = p->lr;
	*beta1t *= p->beta1; /* main.cpp:718-719 */
	*beta2t *= p->beta2;
	float b1 = *beta1t, b2 = *beta2t;
	int bad = 0;
	for( int i = 0; i < n; i++ )
	{
		gs2d_splat* s = &splats[i];
		gs2d_splat_adam* a = &adam[i];
		const gs2d_splat* d = &dsplats[i];
		s->color[0] = adam_optimize( &a->color[0], s->color[0], d->color[0], trainingRate, b1, b2, p );
		s->color[1] = adam_optimize( &a->color[1], s->color[1], d->color[1], trainingRate, b1, b2, p );
		s->color[2] = adam_optimize( &a->color[2], s->color[2], d->color[2], trainingRate, b1, b2, p );
		s->pos[0] = adam_optimize( &a->pos[0], s->pos[0], d->pos[0], trainingRate, b1, b2, p );
		s->pos[1] = adam_optimize( &a->pos[1], s->pos[1], d->pos[1], trainingRate, b1, b2, p );
		s->sx = adam_optimize( &a->sx, s->sx, d->sx, trainingRate, b1, b2, p );
		s->sy = adam_optimize( &a->sy, s->sy, d->sy, trainingRate, b1, b2, p );
		s->rot = adam_optimize( &a->rot, s->rot, d->rot, trainingRate, b1, b2, p );
		if( p->optimize_opacity )
			s->opacity = adam_optimize( &a->opacity, s->opacity, d->opacity, trainingRate, b1, b2, p );

		s->pos[0] = clampf( s->pos[0], 0.0f, (float)W - 1 );
		s->pos[1] = clampf( s->pos[1], 0.0f, (float)H - 1 );
		s->sx = clampf( s->sx, p->s_min, p->s_max );
		s->sy = clampf( s->sy, p->s_min, p->s_max );
		s->color[0] = clampf( s->color[0], 0.0f, 1.0f );
		s->color[1] = clampf( s->color[1], 0.0f, 1.0f );
		s->color[2] = clampf( s->color[2], 0.0f, 1.0f );
		s->opacity = clampf( s->opacity, p->o_min, p->o_max );
	}
	for( int i = 0; i < n; i++ ) /* main.cpp:752-785: pos.y and opacity are not checked */
	{
		const gs2d_splat* s = &splats[i];
		if( !isfinite( s->color[0] ) || !isfinite( s->color[1] ) || !isfinite( s->color[2] ) || !isfinite( s->sx ) ||
		    !isfinite( s->sy ) || !isfinite( s->rot ) || !isfinite( s->pos[0] ) )
			bad = 1;
	}
	return bad;
}

double oracle_mse( int W, int H, const float* image, const float* target )
{
	double mse = 0.0;
	for( int y = 0; y < H; y++ )
		for( int x = 0; x < W; x++ )
		{
			size_t px = (size_t)y * W + x;
			float dx = ( image[4 * px + 0] - target[4 * px + 0] ) * 255.0f;
			float dy = ( image[4 * px + 1] - target[4 * px + 1] ) * 255.0f;
			float dz = ( image[4 * px + 2] - target[4 * px + 2] ) * 255.0f;
			mse += dx * dx + dy * dy + dz * dz; /* lengthSquared, main.cpp:131-134: float, then += into double */
		}
	mse /= ( H * W * 3 );
	return mse;
}

double oracle_step( int W, int H, int n, gs2d_splat* splats, gs2d_splat_adam* adam, float* beta1t, float* beta2t,
                    const float* target, const gs2d_params* p, float* image, float* scratch, gs2d_splat* dsplats,
                    int* nonfinite, uint64_t counts[2] )
{
	gs2d_splat* ds = dsplats ? dsplats : (gs2d_splat*)malloc( sizeof( gs2d_splat ) * (size_t)n );
	oracle_forward( W, H, n, splats, p, image, counts );
	oracle_backward( W, H, n, splats, p, image, target, scratch, ds );
	int bad = oracle_adam( W, H, n, splats, adam, ds, beta1t, beta2t, p );
	if( nonfinite )
		*nonfinite = bad;
	if( !dsplats )
		free( ds );
	return oracle_mse( W, H, image, target );
}
