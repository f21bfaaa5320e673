// j2k::InputFile / j2k::OutputFile over stdio FILE* -- the Linux counterpart of PlatformInputFile / PlatformOutputFile
// (src/common/j2k_platform_io.cpp, Mac Carbon and Win32 only).  Interface: src/common/j2k_io.h:35-79.
#ifndef J2K_POSIX_IO_H
#define J2K_POSIX_IO_H

#include <cstddef>
#include <cstdio>
#include "j2k_io.h"

namespace j2k
{

class PosixInputFile : public InputFile
{
  public:
	PosixInputFile(const char *path);
	virtual ~PosixInputFile();

	virtual ReadFlags Flags() const { return J2K_READ_SEEKABLE; }
	virtual size_t FileSize();
	virtual size_t Read(void *buf, size_t num_bytes);
	virtual bool Seek(size_t position);
	virtual size_t Tell();

  private:
	FILE *_fp;
};

class PosixOutputFile : public OutputFile
{
  public:
	PosixOutputFile(const char *path);
	virtual ~PosixOutputFile();

	virtual WriteFlags Flags() const { return (J2K_WRITE_SEEKABLE | J2K_WRITE_READABLE); }
	virtual size_t Read(void *buf, size_t num_bytes);
	virtual size_t Write(const void *buf, size_t num_bytes);
	virtual bool Seek(size_t position);
	virtual size_t Tell();

  private:
	FILE *_fp;
};

}; // namespace j2k

#endif // J2K_POSIX_IO_H
