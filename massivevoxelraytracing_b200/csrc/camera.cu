// camera.cu -- host-side camera setup (no device code).  CameraPinhole::initFromPerspective (renderCommon.hpp:21-35)
// fed by glm::lookAt + glm::perspectiveFov as pr::GetCameraMatrix does (libs/prlib/src/pr.cpp:1104-1112,1154-1156).
// glm is restated with its operation order (dot = x+y+z of the products, normalize = v * (1/sqrt(dot)), RH, [-1,1] depth).
#include <cmath>

#include "internal.cuh"

namespace
{
struct F3
{
	float x, y, z;
};
inline F3 sub( F3 a, F3 b ) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline float gdot( F3 a, F3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline F3 gcross( F3 x, F3 y ) { return { x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y }; }
inline F3 gnormalize( F3 v )
{
	float inv = 1.0f / sqrtf( gdot( v, v ) );
	return { v.x * inv, v.y * inv, v.z * inv };
}
} // namespace

extern "C" {

int hako_camera_from_matrices( const float view[16], const float proj[16], float focus, float lensR, hako_camera* out )
{
	if( !view || !proj || !out )
	{
		hako::setError( "hako_camera_from_matrices: null argument" );
		return HAKO_E_INVALID;
	}
	// column-major: M[c][r] = m[c * 4 + r].  vT = transpose( mat3( view ) ) => vT[c] = row c of view
	F3 vT0 = { view[0], view[4], view[8] };
	F3 vT1 = { view[1], view[5], view[9] };
	F3 vT2 = { view[2], view[6], view[10] };
	out->front[0] = -vT2.x, out->front[1] = -vT2.y, out->front[2] = -vT2.z;
	out->up[0] = vT1.x, out->up[1] = vT1.y, out->up[2] = vT1.z;
	out->right[0] = vT0.x, out->right[1] = vT0.y, out->right[2] = vT0.z;
	F3 v = { view[12], view[13], view[14] };
	F3 m = { vT0.x * v.x + vT1.x * v.y + vT2.x * v.z, vT0.y * v.x + vT1.y * v.y + vT2.y * v.z, vT0.z * v.x + vT1.z * v.y + vT2.z * v.z };
	out->o[0] = -m.x, out->o[1] = -m.y, out->o[2] = -m.z;
	out->tanHthetaY = 1.0f / proj[5];
	out->lensR = lensR;
	out->focus = focus;
	return HAKO_OK;
}

int hako_camera_lookat( const float eye[3], const float center[3], const float up[3], float fovy, int width, int height,
						float zNear, float zFar, float focus, float lensR, hako_camera* out )
{
	if( !eye || !center || !up || !out || width <= 0 || height <= 0 )
	{
		hako::setError( "hako_camera_lookat: bad argument" );
		return HAKO_E_INVALID;
	}
	F3 e = { eye[0], eye[1], eye[2] }, c = { center[0], center[1], center[2] }, u0 = { up[0], up[1], up[2] };
	F3 f = gnormalize( sub( c, e ) );
	F3 s = gnormalize( gcross( f, u0 ) );
	F3 u = gcross( s, f );
	float view[16] = { s.x, u.x, -f.x, 0.0f, s.y, u.y, -f.y, 0.0f, s.z, u.z, -f.z, 0.0f, -gdot( s, e ), -gdot( u, e ), gdot( f, e ), 1.0f };
	// glm::perspectiveFovRH_NO
	float h = cosf( 0.5f * fovy ) / sinf( 0.5f * fovy );
	float w = h * (float)height / (float)width;
	float proj[16] = { w, 0, 0, 0, 0, h, 0, 0, 0, 0, -( zFar + zNear ) / ( zFar - zNear ), -1.0f, 0, 0, -( 2.0f * zFar * zNear ) / ( zFar - zNear ), 0 };
	return hako_camera_from_matrices( view, proj, focus, lensR, out );
}
}
