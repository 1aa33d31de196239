/* oracle/ref_stubs/DirectXCollision.h — TEST INFRASTRUCTURE.  The one thing RayTracer::createOctTree takes from DirectXMath:
 * BoundingBox::GetCorners = Center + Extents * g_BoxOffset[i] (DirectXCollision.inl: XMVectorMultiplyAdd(extents, offset,
 * center), here as a multiply followed by an add; the offsets are +-1, so the products are exact). */
#pragma once
#include <cstddef>
namespace DirectX {
struct XMFLOAT3 { float x, y, z; XMFLOAT3() = default; XMFLOAT3(float a, float b, float c) : x(a), y(b), z(c) {} };
struct BoundingBox {
    static constexpr size_t CORNER_COUNT = 8;
    XMFLOAT3 Center{0.f, 0.f, 0.f};
    XMFLOAT3 Extents{1.f, 1.f, 1.f};
    void GetCorners(XMFLOAT3* corners) const
    {
        static const float off[8][3] = {{-1, -1, 1}, {1, -1, 1}, {1, 1, 1}, {-1, 1, 1}, {-1, -1, -1}, {1, -1, -1}, {1, 1, -1}, {-1, 1, -1}};
        for (int i = 0; i < 8; i++) {
            volatile float px = Extents.x * off[i][0], py = Extents.y * off[i][1], pz = Extents.z * off[i][2];
            corners[i] = XMFLOAT3(px + Center.x, py + Center.y, pz + Center.z);
        }
    }
};
}
