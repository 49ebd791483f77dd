// host_io.h -- single-process stand-ins for MPI::Communicator (MPIInterface.h:108-431) and
// MPI::LogFile (MPIFile.h:23-56).  One process drives one GPU; rank/size follow torchrun's RANK/WORLD_SIZE.
#pragma once
#include <cstdio>
#include <memory>
#include <string>

namespace pgt_host {

class Communicator {
public:
    Communicator();
    int rank() const { return m_rank; }
    int size() const { return m_size; }
    void barrier() const;
    void abort(int code) const;

private:
    int m_rank = 0, m_size = 1;
};

class LogFile {
public:
    LogFile(std::shared_ptr<Communicator> comm, const std::string& filename);
    ~LogFile();
    LogFile(const LogFile&) = delete;
    LogFile& operator=(const LogFile&) = delete;
    void write_shared(const std::string& s);
    void flush();

private:
    std::FILE* m_file = nullptr;
    std::shared_ptr<Communicator> m_comm;
};

}  // namespace pgt_host
