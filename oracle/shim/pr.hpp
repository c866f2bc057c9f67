// oracle/shim/pr.hpp — TEST INFRASTRUCTURE, not product code.
//
// Headless stand-in for the reference's prlib umbrella header (libs/prlib/src/pr.hpp) so that the
// reference's own main.cpp compiles UNMODIFIED, where it lies under /root/reference, with plain g++
// and no graphics stack ("oracle level 0").  Only the symbols main.cpp touches exist.  Everything
// that draws is a no-op; the three image facts that feed the hot path are real:
//   * Image2DRGBA8::load            = stbi_load(..., 4)          (prp.cpp:179-193)
//   * Image2DRGBA32 container       = row-major vec4, y*w+x      (prp.cpp:347-352, 751-756)
//   * Image2DRGBA8_to_Image2DRGBA32 = vec4(u8)/vec4(255.0f)      (prp.cpp:796-807)
// The reference's own glm and stb_image are used (include paths point into /root/reference).
//
// Environment knobs (read in shim.cpp):
//   PR_DATA   directory holding the JPEGs               (default ".")
//   PR_IMAGE  file name substituted for whatever main.cpp asks to load (config C2)
//   PR_ITERS  number of loop iterations before NextFrame() reports "window closed" (default 20)
//   PR_DUMP   directory; every tex0 upload is written there as itr%05d.rgba32f (raw floats)
//   PR_TARGET file; the decoded u8 RGBA target is written there (w,h as two int32 then bytes)
#pragma once

#include <cmath>
#include <math.h> // libstdc++ wrapper: pulls the float overloads (sqrt(float) ...) into the global namespace, as MSVC does
#include <type_traits>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <functional>

#include "glm/glm.hpp"
#include "glm/ext.hpp"

// main.cpp spells these std::sqrtf / std::cosf / std::sinf (MSVC has them there; libstdc++ does not)
// and calls isfinite unqualified.
namespace std
{
using ::cosf;
using ::sinf;
using ::sqrtf;
} // namespace std
using std::isfinite;
// main.cpp:155 calls unqualified sqrt(float): it must resolve to the float overload (as under MSVC), not ::sqrt(double).
static_assert( std::is_same<decltype( sqrt( 1.0f ) ), float>::value, "sqrt(float) must stay float" );

namespace pr
{
enum class Result
{
	Sucess,
	Failure
};

struct Config
{
	int ScreenWidth = 0;
	int ScreenHeight = 0;
	int SwapInterval = 0;
	int NumSamples = 0;
};

struct Camera3D
{
	glm::vec3 origin = { 0, 0, 1 };
	glm::vec3 lookat = { 0, 0, 0 };
	glm::vec3 up = { 0, 1, 0 };
	float fovy = 0.7f;
	float zNear = 0.1f;
	float zFar = 1000.0f;
};

enum class GridAxis
{
	XY,
	XZ,
	YZ
};
enum class PrimitiveMode
{
	Points,
	Lines,
	LineStrip
};

class Image2DRGBA8
{
public:
	using PixelType = glm::u8vec4;
	Result load( const char* filename );
	int width() const { return _width; }
	int height() const { return _height; }
	const PixelType& operator()( int x, int y ) const { return _values[y * _width + x]; }
	PixelType& operator()( int x, int y ) { return _values[y * _width + x]; }
	const PixelType* data() const { return _values.data(); }

private:
	int _width = 0;
	int _height = 0;
	std::vector<PixelType> _values;
};

class Image2DRGBA32
{
public:
	using PixelType = glm::vec4;
	void allocate( int w, int h )
	{
		_width = w;
		_height = h;
		_values.clear();
		_values.resize( (size_t)_width * _height, glm::vec4( 0.0f, 0.0f, 0.0f, 1.0f ) );
	}
	PixelType* data() { return _values.data(); }
	const PixelType* data() const { return _values.data(); }
	PixelType& operator()( int x, int y ) { return _values[y * _width + x]; }
	const PixelType& operator()( int x, int y ) const { return _values[y * _width + x]; }
	int width() const { return _width; }
	int height() const { return _height; }

private:
	int _width = 0;
	int _height = 0;
	std::vector<PixelType> _values;
};

Image2DRGBA32 Image2DRGBA8_to_Image2DRGBA32( const Image2DRGBA8& src );

class ITexture
{
public:
	void upload( const Image2DRGBA32& image ); // dumps when PR_DUMP is set (shim.cpp)
	int width() const { return _w; }
	int height() const { return _h; }
	int _w = 0, _h = 0, _id = 0, _uploads = 0;
};
ITexture* CreateTexture();

class CircleGenerator
{
public:
	CircleGenerator( float stepThetaT ) : _sd( std::sin( stepThetaT ) ), _cd( std::cos( stepThetaT ) ) {}
	float sin() const { return _s; }
	float cos() const { return _c; }
	void step()
	{
		float ns = _s * _cd + _c * _sd;
		float nc = _c * _cd - _s * _sd;
		_s = ns;
		_c = nc;
	}

private:
	float _sd, _cd, _s = 0.0f, _c = 1.0f;
};

// ---- window / frame loop: headless -------------------------------------------------------------
void SetDataDir( std::string dir );
std::string ExecutableDir();
void Initialize( Config config );
void CleanUp();
bool NextFrame(); // true ("closed") once PR_ITERS loop bodies have run
double GetElapsedTime();
double GetFrameRate();
inline bool IsImGuiUsingMouse() { return true; } // camera update is skipped
inline void UpdateCameraBlenderLike( Camera3D* ) {}
inline void ClearBackground( float, float, float, float ) {}
inline void BeginCamera( Camera3D ) {}
inline void EndCamera() {}
inline void PushGraphicState() {}
inline void PopGraphicState() {}
inline void DrawGrid( GridAxis, float, int, glm::u8vec3 ) {}
inline void DrawXYZAxis( float ) {}
inline void DrawCube( glm::vec3, glm::vec3, glm::u8vec3 ) {}
inline void PrimBegin( PrimitiveMode, float = 1.0f ) {}
// recorded (frame, position, colour) when PR_PRIMS names a file: the overlay main() draws per splat (main.cpp:442-476),
// the golden data of gs2d_overlay_geometry (tests/golden/make_overlay_golden.py)
void PrimVertex( glm::vec3 p, glm::u8vec3 c );
inline void PrimEnd() {}
inline void DrawText( glm::vec3, std::string, float = 16.0f ) {}
inline void BeginImGui() {}
inline void EndImGui() {}
} // namespace pr

struct ImVec2
{
	float x, y;
	ImVec2( float x_ = 0, float y_ = 0 ) : x( x_ ), y( y_ ) {}
};
enum
{
	ImGuiCond_Once = 2
};
namespace ImGui
{
inline void SetNextWindowPos( ImVec2, int = 0 ) {}
inline void SetNextWindowSize( ImVec2, int = 0 ) {}
inline bool Begin( const char* ) { return true; }
inline void End() {}
inline void Text( const char*, ... ) {}
inline bool InputInt( const char*, int* ) { return false; }
inline bool Checkbox( const char*, bool* ) { return false; }
inline bool Button( const char* ) { return false; }
inline void Image( void*, ImVec2 ) {}
} // namespace ImGui
