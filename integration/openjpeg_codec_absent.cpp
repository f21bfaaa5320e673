// The reference's codec registry constructs an OpenJPEGCodec (src/common/j2k_codec.cpp:508-519), whose implementation
// (src/common/j2k_openjpeg_codec.cpp) needs ext/openjpeg -- an EMPTY directory in the reference snapshot.  So that
// the reference's unchanged j2k_codec.cpp links in the test harness, this file supplies the four virtuals of that class;
// each reports that the library is not part of this build.  In the real plug-in this file does not exist.
#include "j2k_openjpeg_codec.h"
#include "j2k_exception.h"

namespace j2k
{

bool OpenJPEGCodec::Verify(InputFile &file) { return Codec::Verify(file); }
void OpenJPEGCodec::GetFileInfo(InputFile &, FileInfo &) { throw Exception("OpenJPEG is not linked into this build"); }
void OpenJPEGCodec::ReadFile(InputFile &, const Buffer &, unsigned int, Progress *) { throw Exception("OpenJPEG is not linked into this build"); }
void OpenJPEGCodec::WriteFile(OutputFile &, const FileInfo &, const Buffer &, Progress *) { throw Exception("OpenJPEG is not linked into this build"); }

}; // namespace j2k
