package main

// Same entry point as backend/go/stores/main.go:18-26: --addr, then the stock gRPC shell.

import (
	"flag"
	"os"

	grpc "github.com/mudler/LocalAI/pkg/grpc"
	"github.com/rs/zerolog"
	"github.com/rs/zerolog/log"
)

var (
	addr = flag.String("addr", "localhost:50051", "the address to connect to")
)

func main() {
	log.Logger = log.Output(zerolog.ConsoleWriter{Out: os.Stderr})
	flag.Parse()
	if err := grpc.StartServer(*addr, NewStore()); err != nil {
		panic(err)
	}
}
