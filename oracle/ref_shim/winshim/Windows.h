// TEST INFRASTRUCTURE. Linux stand-in for the handful of <Windows.h> names utils/image/DDSLoader.cpp and
// DDSLoaderHelpers.h use (types, SAL annotations, HRESULT helpers), so that the reference's DDS parser compiles
// from where it lies under /root/reference and pins csrc/dds.cpp's ParseDds (oracle/build_ref.sh).
#pragma once
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <optional>
#include <vector>
typedef int32_t HRESULT;  // Windows LONG is 32 bits: failure codes must come out negative
typedef unsigned int UINT;
#define S_OK ((HRESULT)0)
#define E_FAIL ((HRESULT)0x80004005u)
#define E_POINTER ((HRESULT)0x80004003u)
#define E_INVALIDARG ((HRESULT)0x80070057u)
#define E_UNEXPECTED ((HRESULT)0x8000FFFFu)
#define ERROR_ARITHMETIC_OVERFLOW 534L
#define ERROR_HANDLE_EOF 38L
#define ERROR_NOT_SUPPORTED 50L
#define HRESULT_FROM_WIN32(x) ((HRESULT)(0x80070000u | ((uint32_t)(x) & 0xFFFFu)))
#define SUCCEEDED(hr) (((HRESULT)(hr)) >= 0)
#define FAILED(hr) (((HRESULT)(hr)) < 0)
#define _In_
#define _In_opt_
#define _Out_
#define _Out_opt_
#define _Inout_
#define _In_reads_(x)
#define _In_reads_bytes_(x)
#define _Outptr_opt_
#define _Use_decl_annotations_
// auto_delete_file (DDSLoaderHelpers.h:790-815, screen-grab helper, never instantiated here)
typedef void* HANDLE;
#define TRUE 1
struct FILE_DISPOSITION_INFO { int DeleteFile; };
enum { FileDispositionInfo = 4 };
inline int SetFileInformationByHandle(HANDLE, int, void*, unsigned) { return 0; }
