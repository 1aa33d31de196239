/* oracle/ref_stubs/d3d11.h — TEST INFRASTRUCTURE.  Stand-in for <d3d11.h> so that the reference's RayTracer.h / DX11.h /
 * GPUStorage.h / RenderTexture.h can be PARSED on Linux when oracle/Makefile compiles RayTracer::createOctTree /
 * loadOctTree / refitOctTreeNode (ref_octree_shim.cpp).  Opaque types and no-op methods only: nothing of D3D11 runs here. */
#pragma once
#include <cstddef>
#include <cstdint>
typedef long HRESULT;
typedef void* HWND;
typedef unsigned int UINT;
#define FAILED(hr) (((HRESULT)(hr)) < 0)
#define SUCCEEDED(hr) (((HRESULT)(hr)) >= 0)
enum DXGI_FORMAT { DXGI_FORMAT_UNKNOWN = 0, DXGI_FORMAT_R32G32B32A32_FLOAT = 2, DXGI_FORMAT_R8G8B8A8_UNORM = 28, DXGI_FORMAT_R8_UNORM = 61 };
enum D3D11_MAP { D3D11_MAP_WRITE_DISCARD = 4 };
struct D3D11_MAPPED_SUBRESOURCE { void* pData; UINT RowPitch; UINT DepthPitch; };
struct ID3D11Resource {};
struct ID3D11Buffer : ID3D11Resource {};
struct DXGI_SAMPLE_DESC { UINT Count, Quality; };
struct D3D11_TEXTURE2D_DESC { UINT Width, Height, MipLevels, ArraySize; DXGI_FORMAT Format; DXGI_SAMPLE_DESC SampleDesc; UINT Usage, BindFlags, CPUAccessFlags, MiscFlags; };
struct D3D11_SUBRESOURCE_DATA { const void* pSysMem; UINT SysMemPitch, SysMemSlicePitch; };
enum { D3D11_BIND_SHADER_RESOURCE = 0x8, D3D11_RESOURCE_MISC_TEXTURECUBE = 0x4 };
struct ID3D11Texture2D : ID3D11Resource { void GetDesc(D3D11_TEXTURE2D_DESC* d) { *d = D3D11_TEXTURE2D_DESC(); } void Release() {} };
struct ID3D11ShaderResourceView {};
struct ID3D11UnorderedAccessView {};
struct ID3D11RenderTargetView {};
struct ID3D11DepthStencilView {};
struct ID3D11ComputeShader {};
/* CreateTexture2D keeps what the caller hands over as initial data (RayTracer::loadEnvironmentMap's six faces: the sink is set
 * by ref_octree_shim.cpp); nothing is created */
struct OkrtTextureSink { void (*take)(const D3D11_TEXTURE2D_DESC*, const D3D11_SUBRESOURCE_DATA*); };
extern OkrtTextureSink g_okrtTextureSink;
struct ID3D11Device {
    HRESULT CreateTexture2D(const D3D11_TEXTURE2D_DESC* desc, const D3D11_SUBRESOURCE_DATA* data, ID3D11Texture2D** out)
    {
        static ID3D11Texture2D texture;
        if (g_okrtTextureSink.take) g_okrtTextureSink.take(desc, data);
        *out = &texture;
        return 0;
    }
    HRESULT CreateShaderResourceView(ID3D11Resource*, const void*, ID3D11ShaderResourceView** out)
    {
        static ID3D11ShaderResourceView view;
        *out = &view;
        return 0;
    }
};
struct IDXGISwapChain {};
struct ID3D11DeviceContext {
    HRESULT Map(ID3D11Resource*, UINT, D3D11_MAP, UINT, D3D11_MAPPED_SUBRESOURCE*) { return -1; }
    void Unmap(ID3D11Resource*, UINT) {}
    void ClearUnorderedAccessViewFloat(ID3D11UnorderedAccessView*, const float[4]) {}
};
