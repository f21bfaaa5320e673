// TEST HARNESS (not part of the product): the reference's UNCHANGED src/common layer (j2k_codec.cpp, j2k_rgba_file.cpp,
// j2k_exception.cpp, j2k_io.cpp; Little-CMS beside them) compiled for Linux, with j2k::CudaCodec handed to
// RGBAoutputFile / RGBAinputFile the way the After Effects glue hands them a codec (src/aftereffects/j2k.cpp:422-501,
// 622-868), and POSIX files underneath.  Built by oracle/Makefile into oracle/_ref/libj2kplugin.so; driven through ctypes by
// tests/test_integration_plugin.py.  Also exposes the reference's own Codec::CopyBuffer as the oracle for row a1.
#include "j2k_cuda_codec.h"
#include "j2k_rgba_file.h"
#include "j2k_exception.h"
#include "posix_io.h"

#include <cstring>
#include <string>

using namespace j2k;

static void SetErr(char *err, int errlen, const char *what)
{
	if(err != NULL && errlen > 0) { strncpy(err, what, errlen - 1); err[errlen - 1] = 0; }
}

// an After Effects world: interleaved A, R, G, B samples of 1 or 2 bytes (WorldToBuffer, src/aftereffects/j2k.cpp:324-362)
static RGBAbuffer WorldToBuffer(void *data, int width, int height, int rowbytes, int sampleBytes)
{
	RGBAbuffer b;
	Channel *channels[4] = { &b.a, &b.r, &b.g, &b.b };
	for(int i=0; i < 4; i++)
	{
		Channel &c = *channels[i];
		c.width = width; c.height = height;
		c.sampleType = (sampleBytes == 2 ? USHORT : UCHAR);
		c.depth = static_cast<unsigned char>(sampleBytes * 8);
		c.sgnd = false;
		c.buf = (unsigned char *)data + (i * sampleBytes);
		c.colbytes = 4 * sampleBytes;
		c.rowbytes = rowbytes;
	}
	return b;
}

extern "C" {

// ARGB world -> RGBAoutputFile::WriteFile -> CudaCodec -> file.  method: 0 LOSSLESS 1 SIZE 2 QUALITY 3 CINEMA (j2k_codec.h)
int itest_write(const char *path, void *argb, int width, int height, int rowbytes, int sampleBytes, int channels, int fileDepth,
				int format, int method, int layers, int tileSize, int fileSizeKiB, int quality, int reversible, int dci4k,
				int asWritten, char *err, int errlen)
{
	try
	{
		CudaCodec::SetMode(asWritten ? CudaCodec::AS_WRITTEN : CudaCodec::FULL);
		static CudaCodec codec;
		FileInfo info;
		info.width = width; info.height = height;
		info.channels = static_cast<unsigned char>(channels);
		info.depth = static_cast<unsigned char>(fileDepth);
		info.format = static_cast<Format>(format);
		info.alpha = (channels == 4 ? STRAIGHT : NO_ALPHA);
		info.colorSpace = sRGB;
		info.settings.method = static_cast<CompressionMethod>(method);
		info.settings.layers = static_cast<unsigned char>(layers);
		info.settings.tileSize = static_cast<unsigned short>(tileSize);
		info.settings.fileSize = fileSizeKiB;
		info.settings.quality = static_cast<unsigned char>(quality);
		info.settings.reversible = (reversible != 0);
		info.settings.order = LRCP;
		info.settings.dciProfile = (dci4k ? DCI_4K : DCI_2K);
		RGBAbuffer buffer = WorldToBuffer(argb, width, height, rowbytes, sampleBytes);
		PosixOutputFile file(path);
		RGBAoutputFile out(file, info, &codec);
		out.WriteFile(buffer);
		return 0;
	}
	catch(std::exception &e) { SetErr(err, errlen, e.what()); return 1; }
	catch(...) { SetErr(err, errlen, "unknown exception"); return 2; }
}

// file -> RGBAinputFile (GetFileInfo in its constructor) -> ReadFile -> ARGB world of width/subsample x height/subsample.
// info_out: width, height, channels, depth, format, alpha, colorSpace, reversible
int itest_read(const char *path, void *argb, int width, int height, int rowbytes, int sampleBytes, int subsample, int *info_out,
			   char *err, int errlen)
{
	try
	{
		static CudaCodec codec;
		PosixInputFile file(path);
		RGBAinputFile in(file, &codec);
		const FileInfo &info = in.GetFileInfo();
		if(info_out != NULL)
		{
			info_out[0] = info.width; info_out[1] = info.height; info_out[2] = info.channels; info_out[3] = info.depth;
			info_out[4] = info.format; info_out[5] = info.alpha; info_out[6] = info.colorSpace; info_out[7] = info.settings.reversible;
		}
		if(argb != NULL)
		{
			RGBAbuffer buffer = WorldToBuffer(argb, width, height, rowbytes, sampleBytes);
			in.ReadFile(buffer, subsample);
		}
		return 0;
	}
	catch(std::exception &e) { SetErr(err, errlen, e.what()); return 1; }
	catch(...) { SetErr(err, errlen, "unknown exception"); return 2; }
}

// the reference's own Codec::CopyBuffer on one channel (src/common/j2k_codec.cpp:403-427): the oracle of row a1
int itest_copy_channel(void *dst, int dstType, int dstDepth, int dstSgnd, long dstCol, long dstRow,
					   void *src, int srcType, int srcDepth, int srcSgnd, long srcCol, long srcRow, int width, int height)
{
	Buffer d, s;
	d.channels = s.channels = 1;
	Channel &dc = d.channel[0], &sc = s.channel[0];
	dc.width = sc.width = width; dc.height = sc.height = height;
	dc.sampleType = static_cast<SampleType>(dstType); dc.depth = static_cast<unsigned char>(dstDepth); dc.sgnd = (dstSgnd != 0);
	dc.buf = (unsigned char *)dst; dc.colbytes = dstCol; dc.rowbytes = dstRow;
	sc.sampleType = static_cast<SampleType>(srcType); sc.depth = static_cast<unsigned char>(srcDepth); sc.sgnd = (srcSgnd != 0);
	sc.buf = (unsigned char *)src; sc.colbytes = srcCol; sc.rowbytes = srcRow;
	Codec::CopyBuffer(d, s);
	return 0;
}

// the codec registry of the unchanged j2k_codec.cpp still works (its OpenJPEGCodec is the stub of openjpeg_codec_absent.cpp)
const char *itest_default_codec_name(void)
{
	try { return GetDefaultCodec()->Name(); } catch(...) { return "?"; }
}

} // extern "C"
