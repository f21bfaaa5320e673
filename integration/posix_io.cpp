#include "posix_io.h"
#include "j2k_exception.h"

namespace j2k
{

PosixInputFile::PosixInputFile(const char *path) : _fp(fopen(path, "rb"))
{
	if(_fp == NULL)
		throw Exception("Couldn't open file for reading.");
}

PosixInputFile::~PosixInputFile()
{
	if(_fp != NULL)
		fclose(_fp);
}

size_t PosixInputFile::FileSize()
{
	const long here = ftell(_fp);
	fseek(_fp, 0, SEEK_END);
	const long size = ftell(_fp);
	fseek(_fp, here, SEEK_SET);
	return (size < 0 ? 0 : static_cast<size_t>(size));
}

size_t PosixInputFile::Read(void *buf, size_t num_bytes) { return fread(buf, 1, num_bytes, _fp); }
bool PosixInputFile::Seek(size_t position) { return (fseek(_fp, static_cast<long>(position), SEEK_SET) == 0); }   // absolute, as j2k_platform_io.cpp
size_t PosixInputFile::Tell() { const long p = ftell(_fp); return (p < 0 ? 0 : static_cast<size_t>(p)); }


PosixOutputFile::PosixOutputFile(const char *path) : _fp(fopen(path, "w+b"))
{
	if(_fp == NULL)
		throw Exception("Couldn't open file for writing.");
}

PosixOutputFile::~PosixOutputFile()
{
	if(_fp != NULL)
		fclose(_fp);
}

size_t PosixOutputFile::Read(void *buf, size_t num_bytes) { return fread(buf, 1, num_bytes, _fp); }
size_t PosixOutputFile::Write(const void *buf, size_t num_bytes) { return fwrite(buf, 1, num_bytes, _fp); }
bool PosixOutputFile::Seek(size_t position) { return (fseek(_fp, static_cast<long>(position), SEEK_SET) == 0); }
size_t PosixOutputFile::Tell() { const long p = ftell(_fp); return (p < 0 ? 0 : static_cast<size_t>(p)); }

}; // namespace j2k
