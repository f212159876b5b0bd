#!/usr/bin/env bash
# Example mutual-TLS material for the gRPC protocol: a throw-away CA plus one server and one client certificate
# valid for 127.0.0.1 / ::1 / localhost.  Writes ca.crt server.{key,crt} client.{key,crt} next to this script - the
# paths Settings.CA_CRT / SERVER_* / CLIENT_* point at.  (Reference: p2pfl/certificates/gen-certs.sh.)
set -euo pipefail
cd "$(dirname "$0")"
DAYS=${DAYS:-365}
SAN="subjectAltName=IP:127.0.0.1,IP:::1,DNS:localhost"
openssl genrsa -out ca.key 4096 2>/dev/null
openssl req -x509 -new -nodes -key ca.key -sha256 -days "$DAYS" -subj "/CN=myfyp-b200-example-ca" -out ca.crt
for who in server client; do
  openssl genrsa -out "$who.key" 2048 2>/dev/null
  openssl req -new -key "$who.key" -subj "/CN=myfyp-b200-$who" -out "$who.csr"
  printf "%s\nextendedKeyUsage=%s\n" "$SAN" "$([ "$who" = server ] && echo serverAuth || echo clientAuth)" > "$who.ext"
  openssl x509 -req -in "$who.csr" -CA ca.crt -CAkey ca.key -CAcreateserial -days "$DAYS" -sha256 -extfile "$who.ext" -out "$who.crt" 2>/dev/null
  rm -f "$who.csr" "$who.ext"
done
rm -f ca.srl
echo "wrote $(pwd)/{ca.crt,server.key,server.crt,client.key,client.crt}"
