// er_server.cpp — the render server: ElevenRender's main loop (src/main.cpp:190-240) over POSIX sockets.
// Listens on TCP :5557 (`--port N` for tests), sends STATUS "ok" on accept, then serves one session at a
// time: COMMAND messages go to the CommandManager, STATUS "close_session" (or EOF) ends the session.
// `--once` leaves after the first session. There is no CPU fallback: without a compute-capability-10 device
// `--start` fails loudly on stderr and `--get_pass` returns zeros.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "er_commands.hpp"

int main(int argc, char** argv) {
    int port = 5557;
    bool once = false;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--port") && i + 1 < argc) port = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--once")) once = true;
        else {
            fprintf(stderr, "usage: er_server [--port N] [--once]\n");
            return 2;
        }
    }
    signal(SIGPIPE, SIG_IGN);
    const int ls = socket(AF_INET, SOCK_STREAM, 0);
    if (ls < 0) { perror("socket"); return 1; }
    int one = 1;
    setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons((uint16_t)port);
    if (bind(ls, (sockaddr*)&addr, sizeof addr) < 0) { perror("bind"); return 1; }
    if (listen(ls, 1) < 0) { perror("listen"); return 1; }
    socklen_t al = sizeof addr;
    getsockname(ls, (sockaddr*)&addr, &al);
    printf("listening %d\n", (int)ntohs(addr.sin_port));  // `--port 0`: the test reads the port the kernel chose
    fflush(stdout);
    for (;;) {
        erh::TCPInterface tcp;
        tcp.fd = accept(ls, nullptr, nullptr);
        if (tcp.fd < 0) { perror("accept"); return 1; }
        setsockopt(tcp.fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
        {
            erh::CommandManager cm;  // a session owns its scene and its render context (src/main.cpp:206-207)
            cm.im = &tcp;
            tcp.write_message(erh::Message::OK());
            while (!tcp.error) {
                const erh::Message msg = tcp.read_message();
                if (msg.type == erh::Message::Type::COMMAND) {
                    cm.execute(msg);
                } else if (msg.type == erh::Message::Type::STATUS) {
                    if (msg.get_string_data() == "close_session") break;
                    fprintf(stderr, "Message received, expected a command but it's not\n");
                }
            }
        }
        close(tcp.fd);
        if (once) break;
    }
    close(ls);
    return 0;
}
