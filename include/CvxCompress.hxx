// CvxCompress.hxx -- public interface of the B200 drop-in for ChevronETC/CvxCompress.
//
// Declarations (names, argument order and types, virtual destructor, extern "C" block) follow
// the reference's CvxCompress.hxx:19-135 (class) and :139-192 (C functions) so that programs
// written against the reference -- Test_With_Generated_Input.cpp:73-90, Compress_SEAM_Basin.cpp,
// Test_Compression.cpp -- compile and link against this library unchanged.  The implementation
// (cvxcompress_b200/csrc/cvx_class.cpp) forwards to the CUDA path behind include/cvx_gpu.h.
//
// Differences a caller can observe:
//  * `vol` and `compressed` may also be device / managed pointers (then nothing is staged);
//  * `num_threads` is accepted and ignored (no OpenMP; no omp_set_num_threads side effect);
//  * block payloads are always in natural block order (the reference with num_threads == 1);
//  * failures print a message and abort(), where the reference asserts or exits.
#include <stdbool.h>

#ifndef CVX_CVXCOMPRESS_HXX
#define CVX_CVXCOMPRESS_HXX

#ifdef __cplusplus
extern "C" {
#endif

#ifdef __cplusplus

class CvxCompress
{
public:
	CvxCompress();
	virtual ~CvxCompress();

	// nx is the fastest axis, nz the slowest.  `scale` is the relative quantisation threshold
	// (1e-2 .. 1e-5 for seismic wavefields; larger = lossier).  Returns the compression ratio
	// 4*nx*ny*nz / compressed_length and sets compressed_length (bytes).
	float Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
	               bool use_local_RMS, unsigned int* compressed, int num_threads, long& compressed_length);
	float Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
	               bool use_local_RMS, unsigned int* compressed, long& compressed_length);
	// use_local_RMS = false
	float Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
	               unsigned int* compressed, long& compressed_length);
	float Compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
	               unsigned int* compressed, int num_threads, long& compressed_length);

	// out of place: reads nx,ny,nz from the stream and returns a 64-byte aligned host buffer
	// the caller releases with free()
	float* Decompress(int& nx, int& ny, int& nz, unsigned int* compressed, long compressed_length);
	// in place: nx,ny,nz must match the stream
	void Decompress(float* vol, int nx, int ny, int nz, unsigned int* compressed, int num_threads, long compressed_length);
	void Decompress(float* vol, int nx, int ny, int nz, unsigned int* compressed, long compressed_length);

	bool Is_Valid_Block_Size(int bx, int by, int bz);

	static int Min_BX() {return   8;}
	static int Max_BX() {return 256;}
	static int Min_BY() {return   8;}
	static int Max_BY() {return 256;}
	static int Min_BZ() {return   8;}
	static int Max_BZ() {return 256;}

	// GPU self-test: transform round trips, gather/scatter, RMS, compress/decompress on synthetic
	// volumes.  Returns true when every check passes.
	bool Run_Module_Tests(bool verbose, bool exhaustive_throughput_tests);
};

#endif // __cplusplus

float cvx_compress(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                   unsigned int* compressed, long* compressed_length);

float* cvx_decompress_outofplace(int* nx, int* ny, int* nz, unsigned int* compressed, long compressed_length);

void cvx_decompress_inplace(float* vol, int nx, int ny, int nz, unsigned int* compressed, long compressed_length);

float cvx_compress_th(float scale, float* vol, int nx, int ny, int nz, int bx, int by, int bz,
                      bool use_local_RMS, unsigned int* compressed, int num_threads, long* compressed_length);

void cvx_decompress_inplace_th(float* vol, int nx, int ny, int nz, unsigned int* compressed,
                               int num_threads, long compressed_length);

#ifdef __cplusplus
}
#endif

#endif
