// er_message.hpp — the render-command wire format (src/Managers.h:14,101-144, src/Managers.cpp:6-17,167-177,
// src/TCPInterface.cpp:3-58): a frame is a 1024-byte NUL-padded JSON header
//   {"type":"none|command|status|data","data_format":"none|float3|float4|string|json","data_size":N}
// followed by N payload bytes. Plain POSIX sockets: the protocol, not Boost.Asio, is the contract.
#pragma once
#include <errno.h>
#include <string.h>
#include <sys/socket.h>
#include <sys/types.h>
#include <unistd.h>

#include <memory>
#include <string>
#include <vector>

#include "er_json.hpp"

namespace erh {

constexpr size_t MESSAGE_HEADER_SIZE = 1024;  // src/Managers.h:14

struct Message {
    enum class Type { NONE, COMMAND, STATUS, DATA };
    enum class DataFormat { NONE, FLOAT3, FLOAT4, STRING, JSON };

    Type type = Type::NONE;
    DataFormat data_format = DataFormat::NONE;
    std::vector<char> data;  // payload (the reference mallocs it and never frees it, src/TCPInterface.cpp:46-48)

    static Message status(const std::string& s) {
        Message m;
        m.type = Type::STATUS;
        m.data_format = DataFormat::STRING;
        m.data.assign(s.begin(), s.end());
        return m;
    }
    static Message OK() { return status("ok"); }                       // src/Managers.h:111-118
    static Message CloseSession() { return status("close_session"); }  // src/Managers.h:120-127
    static Message json(const erj::Value& v) {
        Message m;
        m.type = Type::DATA;
        m.data_format = DataFormat::JSON;
        const std::string s = erj::serialize(v);
        m.data.assign(s.begin(), s.end());
        return m;
    }
    static Message command(const std::string& s) {
        Message m = status(s);
        m.type = Type::COMMAND;
        return m;
    }

    static const char* type2str(Type t) {  // src/Managers.cpp:44-63
        switch (t) {
            case Type::NONE: return "none";
            case Type::COMMAND: return "command";
            case Type::STATUS: return "status";
            case Type::DATA: return "data";
        }
        return "und";
    }
    static Type str2type(const std::string& s) {  // src/Managers.cpp:65-82: anything else is NONE
        if (s == "command") return Type::COMMAND;
        if (s == "status") return Type::STATUS;
        if (s == "data") return Type::DATA;
        return Type::NONE;
    }
    static const char* data_format2str(DataFormat f) {  // src/Managers.cpp:84-106
        switch (f) {
            case DataFormat::NONE: return "none";
            case DataFormat::STRING: return "string";
            case DataFormat::JSON: return "json";
            case DataFormat::FLOAT3: return "float3";
            case DataFormat::FLOAT4: return "float4";
        }
        return "und";
    }
    static DataFormat str2data_format(const std::string& s) {  // src/Managers.cpp:108-127
        if (s == "float3") return DataFormat::FLOAT3;
        if (s == "float4") return DataFormat::FLOAT4;
        if (s == "json") return DataFormat::JSON;
        if (s == "string") return DataFormat::STRING;
        return DataFormat::NONE;
    }

    std::string header() const {  // msg2json_header + the NUL padding of write_message
        erj::Value h = erj::Value::make_object();
        h.set("type", erj::Value::of(type2str(type)));
        h.set("data_format", erj::Value::of(data_format2str(data_format)));
        h.set("data_size", erj::Value::of((unsigned long long)data.size()));
        std::string s = erj::serialize(h);
        if (s.size() > MESSAGE_HEADER_SIZE) throw std::runtime_error("TCP header size exceeded");
        s.resize(MESSAGE_HEADER_SIZE, '\0');
        return s;
    }
    // json2header (src/Managers.cpp:6-17); returns the payload size the header announces
    static size_t parse_header(const char* bytes, Message& m) {
        const std::string s(bytes, strnlen(bytes, MESSAGE_HEADER_SIZE));
        const erj::Value h = erj::parse(s);
        m.type = str2type(h.at("type").as_string());
        m.data_format = str2data_format(h.at("data_format").as_string());
        const long long n = h.at("data_size").as_int64();
        if (n < 0) throw std::runtime_error("negative data_size");
        return (size_t)n;
    }

    std::string get_string_data() const { return std::string(data.begin(), data.end()); }
    erj::Value get_json_data() const { return erj::parse(get_string_data()); }
    const float* get_float_data() const { return reinterpret_cast<const float*>(data.data()); }
    size_t float_count() const { return data.size() / sizeof(float); }
};

// TCPInterface (src/TCPInterface.h:4-16) over a connected socket
class TCPInterface {
   public:
    int fd = -1;
    bool error = false;  // EOF or a socket error: the session loop ends (src/main.cpp:215)

    void write_message(const Message& m) {
        const std::string h = m.header();
        if (!write_all(h.data(), h.size())) return;
        if (!m.data.empty() && m.data_format != Message::DataFormat::NONE) write_all(m.data.data(), m.data.size());
    }
    // payload straight from a caller-owned buffer (a pass is w*h*16 bytes: no copy into a Message)
    void write_data(Message::DataFormat f, const void* p, size_t bytes) {
        Message m;
        m.type = Message::Type::DATA;
        m.data_format = f;
        m.data.resize(0);
        erj::Value h = erj::Value::make_object();
        h.set("type", erj::Value::of("data"));
        h.set("data_format", erj::Value::of(Message::data_format2str(f)));
        h.set("data_size", erj::Value::of((unsigned long long)bytes));
        std::string s = erj::serialize(h);
        s.resize(MESSAGE_HEADER_SIZE, '\0');
        if (!write_all(s.data(), s.size())) return;
        write_all(p, bytes);
    }
    Message read_message() {
        Message m;
        char hdr[MESSAGE_HEADER_SIZE + 1];
        hdr[MESSAGE_HEADER_SIZE] = '\0';
        if (!read_all(hdr, MESSAGE_HEADER_SIZE)) return Message::CloseSession();  // EOF (src/TCPInterface.cpp:33-36)
        try {
            const size_t n = Message::parse_header(hdr, m);
            if (n) {
                m.data.resize(n);
                if (!read_all(m.data.data(), n)) return Message::CloseSession();
            }
        } catch (const std::exception& e) {
            fprintf(stderr, "Error reading message: %s\n", e.what());  // src/TCPInterface.cpp:53-55
            m = Message();
        }
        return m;
    }

   private:
    bool write_all(const void* p, size_t n) {
        const char* c = static_cast<const char*>(p);
        while (n) {
            const ssize_t w = ::send(fd, c, n, MSG_NOSIGNAL);
            if (w <= 0) {
                if (w < 0 && errno == EINTR) continue;
                error = true;
                return false;
            }
            c += w;
            n -= (size_t)w;
        }
        return true;
    }
    bool read_all(void* p, size_t n) {
        char* c = static_cast<char*>(p);
        while (n) {
            const ssize_t r = ::recv(fd, c, n, 0);
            if (r <= 0) {
                if (r < 0 && errno == EINTR) continue;
                error = true;
                return false;
            }
            c += r;
            n -= (size_t)r;
        }
        return true;
    }
};

}  // namespace erh
