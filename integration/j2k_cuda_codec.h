// j2k::CudaCodec -- the reference-side binding of libj2kcuda: a subclass of the plug-in's abstract
// j2k::Codec (src/common/j2k_codec.h:279-326), written the way OpenJPEGCodec is
// (src/common/j2k_openjpeg_codec.h:33-52).  It lives in the PLUG-IN's tree (src/common/) and is registered
// with one line in CodecContainer::CodecContainer() (src/common/j2k_codec.cpp:508-519):
//     _list.push_back(new CudaCodec);
// Everything above it (RGBAinputFile/RGBAoutputFile, the After Effects glue, dialogs) stays unchanged.
#ifndef J2K_CUDA_CODEC_H
#define J2K_CUDA_CODEC_H

#include "j2k_codec.h"

namespace j2k
{

class CudaCodec : public Codec
{
  public:
	CudaCodec() {}
	virtual ~CudaCodec() {}

	virtual const char * Name() const { return "CUDA"; }          // sorts before "OpenJPEG": becomes the default codec
	virtual const char * FourCharCode() const { return "cuda"; }

	virtual ReadFlags GetReadFlags() { return (J2K_CAN_READ | J2K_CAN_SUBSAMPLE); }
	virtual WriteFlags GetWriteFlags() { return (J2K_CAN_WRITE); }

	virtual void GetFileInfo(InputFile &file, FileInfo &info);
	virtual void ReadFile(InputFile &file, const Buffer &buffer, unsigned int subsample = 1, Progress *progress = NULL);

	virtual void WriteFile(OutputFile &file, const FileInfo &info, const Buffer &buffer, Progress *progress = NULL);

	// AS_WRITTEN: honour exactly the settings OpenJPEGCodec::WriteFile honours (byte-identical output); FULL: all of them.
	enum Mode { AS_WRITTEN = 0, FULL = 1 };
	static void SetMode(Mode m);
	static Mode GetMode();
};

}; // namespace j2k

#endif // J2K_CUDA_CODEC_H
